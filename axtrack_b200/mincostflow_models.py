"""Mirror of the reference's ``transition_model`` (axtrack/mincostflow_models.py:67-119) and of the arc
cost libmot derives from it (libmot/libmot/data_association/offline_mincostflow.py:192-198).

``vis_sim_weight == 0`` (deployed): an A* distance is an integer in [0, max_px] and the time gap one of
``max_num_misses + 1`` values, so the reference's float64 expression is evaluated ONCE per (gap, distance)
on the host with numpy -- the reference's own arithmetic, bit-identical by construction -- and the per-pair
matrices are gathered from that table on the GPU (K3, ``axb_transition_cost_batch``).

``vis_sim_weight != 0``: the O(n_pred * n_cur) Python double loop of ``cv2.compareHist`` calls
(mincostflow_models.py:106-113) is ONE kernel launch (K7, ``axb_vis_sim``: bit-identical to cv2 4.13); the
five float64 operations that follow (:115-118) are the reference's own numpy expression on the host,
because the arc cost ``int(1e5 * cost)`` is discontinuous and numpy's ``log`` is not reproducible bit for
bit on a device.  There is no CPU path for the kernels: without the CUDA library these functions raise.
"""
import numpy as np

MULTI_FACTOR = 1e5          # offline_mincostflow.py:13
NO_ARC = np.iinfo(np.int64).min


def transition_table(max_px_assoc_dist=500, miss_rate=0.6, time_gap=1, vis_sim_weight=0.0):
    """float64 [max_px + 1]: ``transition_model`` cost of every possible A* distance
    (mincostflow_models.py:103-118 with vis_sim = nan_to_num(...) finite and weight 0)."""
    if vis_sim_weight != 0:
        raise ValueError('a cost TABLE only exists for vis_sim_weight == 0; use transition_model(...) (K7) otherwise')
    d = np.arange(int(max_px_assoc_dist) + 1, dtype=np.int64)
    distances = ((d / max_px_assoc_dist) - 1) * -1
    inf_dist = distances == 0
    vis_sim = np.zeros_like(distances)
    costs = -np.log((1 - vis_sim_weight) * distances * (miss_rate ** (time_gap - 1))
                    + vis_sim_weight * vis_sim
                    + 1e-6)
    costs[inf_dist] = np.inf
    return costs


def transition_tables(max_px_assoc_dist=500, miss_rate=0.6, max_num_misses=1, vis_sim_weight=0.0, cost_threshold=None):
    """(cost float64 [max_num_misses + 1, max_px + 1], arc int64 same shape | None).  arc =
    ``int(multi_factor * cost)`` where libmot adds an arc (cost <= cost_threshold,
    offline_mincostflow.py:194-198), ``NO_ARC`` elsewhere."""
    cost = np.stack([transition_table(max_px_assoc_dist, miss_rate, g, vis_sim_weight)
                     for g in range(1, max_num_misses + 2)])
    arc = None
    if cost_threshold is not None:
        arc = np.full(cost.shape, NO_ARC, np.int64)
        for g in range(cost.shape[0]):
            for d in range(cost.shape[1]):
                c = cost[g, d]
                if not (c > cost_threshold):
                    arc[g, d] = int(MULTI_FACTOR * c)
    return cost, arc


def transition_model(**kwargs):
    """Drop-in for ``mincostflow_models.transition_model`` (same keyword arguments, same float64 [N, M]
    result).  Weight 0: the matrix is gathered on the GPU from the per-distance table (K3).  Otherwise the
    visual similarities come from K7 and the cost expression is the reference's (host numpy)."""
    import torch
    from . import ops
    from .ops_pairs import make_pairs
    time_gap = int(kwargs['time_gap'])
    frame_idx = kwargs['frame_idx']
    lbl = f'{kwargs["dataset_name"]}_t:{frame_idx:0>3}-t:{frame_idx - time_gap:0>3}'      # mincostflow_models.py:101
    d = np.ascontiguousarray(kwargs['astar_dists'][lbl])
    max_px = int(kwargs['max_px_assoc_dist'])
    w = kwargs.get('vis_sim_weight', 0.0)
    n, m = d.shape
    if not torch.cuda.is_available():
        raise RuntimeError('axtrack_b200 needs a CUDA device; there is no CPU fallback')
    dev = torch.device('cuda', torch.cuda.current_device())
    pl = torch.from_numpy(d.astype(np.int32).reshape(-1)).to(dev)
    if w == 0:
        tab = transition_table(max_px, kwargs['miss_rate'], time_gap, 0.0)
        pairs, _, _ = make_pairs([(0, 0, n, 0, m, 1)])
        cost, _ = ops.transition_cost_batch(pl, pairs, tab[None, :])
        ops.raise_status(device=dev)
        return cost.cpu().numpy().reshape(n, m)
    as_dev = lambda f: (f if isinstance(f, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(f, dtype=np.float32))).to(dev, torch.float32)
    vis_sim = ops.vis_sim(as_dev(kwargs['predecessor_features']), as_dev(kwargs['features']), pl.view(n, m), max_px)
    ops.raise_status(device=dev)
    vis_sim = vis_sim.cpu().numpy()
    # mincostflow_models.py:103-104,115-118 -- the reference's expression, term by term
    distances = ((kwargs['astar_dists'][lbl] / kwargs['max_px_assoc_dist']) - 1) * -1
    inf_dist = distances == 0
    costs = -np.log((1 - w) * distances * (kwargs['miss_rate'] ** (time_gap - 1))
                    + w * vis_sim
                    + 1e-6)
    costs[inf_dist] = np.inf
    return costs


def feature_model(**kwargs):
    """Drop-in for mincostflow_models.feature_model (mincostflow_models.py:30-64): 180-bin
    histograms of the image crops under the boxes, min-max normalised -> float32 (N, 180, 1);
    ``np.zeros((0,))`` for no boxes, like the reference.  ``image`` may be a numpy array (copied to
    the current CUDA device once per call) or a CUDA tensor; ``boxes`` (N, 4) rows (x, y, w, h)."""
    import torch
    from . import ops
    assert 'image' in kwargs and 'boxes' in kwargs, 'Parameters must contail image and boxes'
    boxes, image = kwargs['boxes'], kwargs['image']
    if len(boxes) == 0:
        return np.zeros((0,))
    dev = image.device if isinstance(image, torch.Tensor) and image.is_cuda else torch.device('cuda', torch.cuda.current_device())
    img = image if isinstance(image, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(image, dtype=np.float32))
    b = torch.from_numpy(np.ascontiguousarray(np.atleast_2d(np.asarray(boxes)), dtype=np.int32))
    feats = ops.feature_hist(img.to(dev, torch.float32), b.to(dev))
    return feats.cpu().numpy()[:, :, None]
