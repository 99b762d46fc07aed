import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, '.')
import bench
from axtrack_b200 import sharding
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE']); local = int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local); dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
wl = dict(bench.WORKLOADS['c2'], T=24)
mask, Y, tt, t0 = bench.make_block(wl)[:4]          # the same 24-frame timelapse on every rank
Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)
ref, _ = sharding.associate_sharded(Yd, td, md)
slabs = sharding.PeerSlabs(dev, cap_elems=1000, cap_pairs=4)     # tiny: forces the collective growth path
for it in range(3):
    got, res = sharding.associate_sharded(Yd, td, md, slabs=slabs)
    torch.cuda.synchronize()
    if rank == 0:
        assert set(got) == set(ref) and len(got) == 2 * 24 - 3, (len(got), len(ref))
        for k in ref:
            assert np.array_equal(got[k][0], ref[k][0]) and np.array_equal(got[k][1], ref[k][1]), k
        print('iter', it, 'peer slabs == nccl gather:', len(got), 'pairs,', sum(v[0].size for v in got.values()), 'entries, cap', slabs.cap_elems, flush=True)
dist.barrier()
dist.destroy_process_group()
