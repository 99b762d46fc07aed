"""Generates tests/golden/heap_tie_cases.npz: pairs where the closed form (oracle.cost_matrix_rule)
differs from the per-pair forked A* -- the heap-tie corner of DESIGN.md section 2 -- with the answer of
the COMPILED reference (oracle/_ref, /root/reference/pyastar2d/src/cpp/astar.cpp unmodified).

    python tests/golden/make_heap_tie_cases.py 100 164          # seeds [100, 164): 1.8 M pairs, ~4 cases
    python tests/golden/make_heap_tie_cases.py 1000 1160        # 4.6 M pairs, 45 cases (all flagged tie regime)

The committed fixture holds the 16 cases with the smallest masks of those two sweeps."""
import sys, numpy as np, pickle, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import oracle
from multiprocessing import Pool

def run(seed):
    rng = np.random.default_rng(seed)
    found = []
    tot = 0
    for it in range(400):
        H, W = int(rng.integers(12, 90)), int(rng.integers(12, 90))
        m = np.zeros((H, W), bool)
        mode = it % 4
        if mode == 3:
            m = rng.random((H, W)) < rng.uniform(0.35, 0.8)
        else:
            for _ in range(int(rng.integers(2, 7))):
                y = int(rng.integers(0, H)); w_ = int(rng.integers(1, 6)); m[y:y + w_, :] = True
                x = int(rng.integers(0, W)); w_ = int(rng.integers(1, 6)); m[:, x:x + w_] = True
            for _ in range(int(rng.integers(0, 8))):
                y, x, g = int(rng.integers(0, H)), int(rng.integers(0, W)), int(rng.integers(1, 5))
                m[y:y + g, x:x + g] = False
        max_px = int(rng.choice([5, 8, 12, 17, 25, 40, 80]))
        fg = np.argwhere(m)
        if len(fg) == 0: continue
        def pick(n):
            p = fg[rng.integers(0, len(fg), n)] + rng.integers(-2, 3, (n, 2)) * (rng.random((n, 1)) < 0.4)
            p[:, 0] = np.clip(p[:, 0], 0, H - 1); p[:, 1] = np.clip(p[:, 1], 0, W - 1)
            return p
        src, dst = pick(6), pick(12)
        a = oracle.cost_matrix_astar(m, src, dst, max_px)
        r = oracle.cost_matrix_rule_ex(m, src, dst, max_px)
        flagged = getattr(run, 'flagged', 0) + int(r[3].sum()); run.flagged = flagged
        tot += a[0].size
        bad = np.argwhere((a[0] != r[0]) | (a[1] != r[1]))
        for i, j in bad:
            found.append(dict(mask=m.copy(), src=src[i].copy(), dst=dst[j].copy(), max_px=max_px,
                              astar=(int(a[0][i, j]), int(a[1][i, j])), rule=(int(r[0][i, j]), int(r[1][i, j])), tie=int(r[3][i, j])))
    return seed, tot, found, getattr(run, 'flagged', 0)

if __name__ == '__main__':
    seeds = range(int(sys.argv[1]), int(sys.argv[2]))
    allf, T = [], 0
    with Pool(8) as p:
        for seed, tot, found, fl in p.imap_unordered(run, seeds):
            T += tot; allf += found; FL = fl
            print(seed, tot, len(found), flush=True)
    print('total pairs', T, 'mismatches', len(allf), 'unflagged mismatches', sum(1 for c in allf if not c['tie']))
    allf.sort(key=lambda c: c['mask'].size)
    out = {}
    for i, c in enumerate(allf[:16]):
        m = c['mask']
        pl, un = oracle.ref_cost_matrix(m, [c['src']], [c['dst']], c['max_px'], threads=False)
        assert (int(pl[0, 0]), int(un[0, 0])) == c['astar']
        out[f'mask_{i}'] = np.packbits(m.astype(np.uint8)); out[f'shape_{i}'] = np.array(m.shape)
        out[f'src_{i}'] = c['src'].astype(np.int32); out[f'dst_{i}'] = c['dst'].astype(np.int32)
        out[f'max_px_{i}'] = np.int32(c['max_px'])
        out[f'ref_{i}'] = np.array([int(pl[0, 0]), int(un[0, 0])], np.int32)
        out[f'rule_{i}'] = np.array(c['rule'], np.int32)
    out['n'] = np.int32(min(len(allf), 16))
    np.savez_compressed(sys.argv[3] if len(sys.argv) > 3 else 'heap_tie_cases.npz', **out)
