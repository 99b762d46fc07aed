"""Arc-list construction for libmot's ``MinCostFlowTracker`` from arrays (second half).

The reference's tracker adds the transition arcs of a frame with a Python double loop over EVERY
(predecessor, detection) pair (libmot/libmot/data_association/offline_mincostflow.py:166-198):
``transition_model(...)``, then ``for i ... for j ...: if cost > cost_threshold: continue;
AddArcWithCapacityAndUnitCost(pred_node + 1, node, 1, int(1e5 * cost))``.  ``make_arc_tracker`` derives a
tracker class whose ``process`` issues exactly the same ``AddArc...`` calls in exactly the same order, but
only for the pairs that HAVE an arc, taken from arrays:

* ``vis_sim_weight == 0``: K3 (``axb_transition_cost_batch``) maps the A* distance matrices of ALL frame
  pairs to arc costs in one launch when the first frame is processed; per (frame, gap) the arcs are the
  entries that are not ``NO_ARC`` (row-major = the loop order of the reference);
* otherwise the cost matrix of the (frame, gap) pair comes from ``transition_model`` of this package (K7).

Nothing of libmot is re-implemented: the node / observation half of ``process`` runs in the parent class
(with a transition model that yields no rows, so its own arc loop adds nothing), and the transition arcs
follow -- the observation arcs of a frame precede its transition arcs in the reference too.
"""
import numpy as np

from . import mincostflow_models as mm


def _no_rows(**kwargs):
    return ()


class ArcTables:
    """arcs(frame_idx, time_gap) -> (i, j, arc) for ``vis_sim_weight == 0``, built with ONE K3 launch."""

    def __init__(self, astar_dists, dataset_name, max_px, miss_rate, max_num_misses, cost_threshold):
        import torch
        from . import ops
        from .ops_pairs import make_pairs
        if not torch.cuda.is_available():
            raise RuntimeError('axtrack_b200 needs a CUDA device; there is no CPU fallback')
        self.name = dataset_name
        keys, entries, flat = [], [], []
        for key, d in astar_dists.items():
            t, tb = int(key.split('_t:')[-1][:3]), int(key.split('-t:')[-1][:3])
            gap = t - tb
            if not 1 <= gap <= max_num_misses + 1:
                continue
            d = np.asarray(d)
            keys.append(key)
            entries.append((0, 0, d.shape[0], 0, d.shape[1] if d.ndim == 2 else 0, gap))
            flat.append(d.astype(np.int32).reshape(-1))
        self.shape = {k: (e[2], e[4]) for k, e in zip(keys, entries)}
        self.off = {}
        self.arc = np.zeros(0, np.int64)
        if keys:
            pairs, total, _ = make_pairs(entries)
            for k, pd in zip(keys, pairs):
                self.off[k] = int(pd['out_off'])
            dev = torch.device('cuda', torch.cuda.current_device())
            cost_tab, arc_tab = mm.transition_tables(max_px, miss_rate, max_num_misses, 0.0, cost_threshold)
            pl = torch.from_numpy(np.concatenate(flat) if total else np.zeros(0, np.int32)).to(dev)
            if total:
                _, arc = ops.transition_cost_batch(pl, pairs, cost_tab, arc_tab)
                ops.raise_status(device=dev)
                self.arc = arc.cpu().numpy()

    def arcs(self, frame_idx, time_gap):
        key = f'{self.name}_t:{frame_idx:0>3}-t:{frame_idx - time_gap:0>3}'          # mincostflow_models.py:101
        n, m = self.shape[key]
        a = self.arc[self.off[key]:self.off[key] + n * m].reshape(n, m)
        i, j = np.nonzero(a != mm.NO_ARC)                                            # row-major: i ascending, then j
        return i, j, a[i, j]


def make_arc_tracker(tracker_cls):
    """-> subclass of libmot's ``MinCostFlowTracker`` (pass the class) that builds its transition arcs from arrays."""

    class ArcTracker(tracker_cls):
        __doc__ = tracker_cls.__doc__

        def process(self, boxes, scores, image=None, features=None, **kwargs):
            slices = [list(x) for x in self.nodes_in_timestep[-(1 + self.max_num_misses):]]
            frame_idx = kwargs['frame_idx']
            real = self.transition_model
            self.transition_model = _no_rows
            try:
                out = super().process(boxes, scores, image=image, features=features, **kwargs)
            finally:
                self.transition_model = real
            node_ids = self.nodes_in_timestep[-1]
            if not node_ids:
                return out
            heads_all = np.asarray(node_ids, np.int64)
            for k, pred in enumerate(slices):                                        # offline_mincostflow.py:168-198
                if not pred:
                    continue
                gap = len(slices) - k
                if self.vis_sim_weight == 0:
                    if getattr(self, '_arc_tables', None) is None:
                        self._arc_tables = ArcTables(self.astar_dists, self.dataset_name, self.max_px_assoc_dist, self.miss_rate,
                                                     self.max_num_misses, self.cost_threshold)
                    i, j, arc = self._arc_tables.arcs(frame_idx, gap)
                else:
                    cur_feats = np.asarray([self.node[x]['feature'] for x in node_ids])
                    pred_feats = np.asarray([self.node[x]['feature'] for x in pred])
                    cost = mm.transition_model(miss_rate=self.miss_rate, time_gap=gap, features=cur_feats,
                                               predecessor_features=pred_feats, frame_idx=frame_idx, astar_dists=self.astar_dists,
                                               vis_sim_weight=self.vis_sim_weight, dataset_name=self.dataset_name,
                                               max_px_assoc_dist=self.max_px_assoc_dist)
                    i, j = np.nonzero(~(cost > self.cost_threshold))                 # `if cost > threshold: continue`
                    arc = (mm.MULTI_FACTOR * cost[i, j]).astype(np.int64)            # int(): truncation toward zero
                tails = np.asarray(pred, np.int64)[i] + 1
                heads = heads_all[j]
                add = self.graph.AddArcWithCapacityAndUnitCost
                for t, h, a in zip(tails.tolist(), heads.tolist(), np.asarray(arc).tolist()):
                    add(t, h, 1, a)
            return out

    ArcTracker.__name__ = 'MinCostFlowTracker'
    return ArcTracker
