"""Device-resident tiling front end of the detector (K6).

``FrameTiler`` is what ``Timelapse.construct_tiles`` (image half, /root/reference/axtrack/
Timelapse.py:508-525, 539-558) and ``Timelapse.get_frametiles_stack`` (:150-157) compute, without the
tiled copy of the whole timelapse and without the per-tile Python loop: the dense X stays in HBM, the
non-empty tile flags come from one launch, and the model input of a BATCH of frames from another, so
``YOLO_AXTrack.detect_axons`` (machinelearning/model.py:119-125) can run one forward per batch instead
of one per frame.  No CPU path: without the CUDA library these calls raise.
"""
import torch

from . import ops


class FrameTiler:
    def __init__(self, X, tilesize=512, temporal_context=2, use_motion_filtered='exclude', timepoints_indices=None):
        """X float32 CUDA [T_raw, C, H, W] (``dataset.X.to_dense()``: channel 0 image, 1 / 2 motion,
        Timelapse.py:426-433).  ``timepoints_indices[t]`` = raw frame of timepoint t (Timelapse.py:113);
        default: the frames that have a full temporal context."""
        self.X = X.contiguous()
        self.tilesize, self.temporal_context = int(tilesize), int(temporal_context)
        self.use_motion_filtered = use_motion_filtered
        T = X.shape[0]
        self.timepoints_indices = (list(range(self.temporal_context, T - self.temporal_context))
                                   if timepoints_indices is None else [int(i) for i in timepoints_indices])
        #: bool [ytiles, xtiles, T_raw] = tile_info[..., 0] (Timelapse.py:551)
        self.nonempty = ops.tile_nonempty(self.X, self.tilesize)
        #: int32 [ntiles, 2] (tile_y, tile_x): the tile axis of X_tiled (Timelapse.py:558), K1's tile_yx
        self.tile_yx = ops.compact_tiles(self.nonempty)

    @property
    def ntiles(self):
        return int(self.tile_yx.shape[0])

    def __len__(self):
        return len(self.timepoints_indices)

    def get_frametiles_batch(self, timepoints):
        """-> float32 CUDA [B, ntiles, nchnls, ts, ts] for the timepoints given."""
        frames = torch.tensor([self.timepoints_indices[int(t)] for t in timepoints], dtype=torch.int32,
                              device=self.X.device)
        out = ops.gather_frametiles(self.X, self.tile_yx, frames, self.tilesize, self.temporal_context,
                                    self.use_motion_filtered)
        return out

    def get_frametiles_stack(self, timepoint):
        """The X of Timelapse.get_frametiles_stack(timepoint): [ntiles, nchnls, ts, ts]."""
        return self.get_frametiles_batch([timepoint])[0]

    def stitched_frame(self, timepoint):
        """AxonDetections.get_frame_and_truedets(timepoint)[0]: float32 CUDA [nch, H, W] (no tile round trip)."""
        return ops.stitched_frame(self.X, self.nonempty, self.timepoints_indices[int(timepoint)], self.tilesize,
                                  self.use_motion_filtered)

    def detect_frames(self, model, timepoints, frames_per_forward=8):
        """Head tensors [B, ntiles, Sx, Sy, 3] of the timepoints: ``model.detect_axons`` on
        ``frames_per_forward`` frames per forward (all their tiles as one batch)."""
        heads = []
        timepoints = list(timepoints)
        for i in range(0, len(timepoints), int(frames_per_forward)):
            chunk = timepoints[i:i + int(frames_per_forward)]
            Xb = self.get_frametiles_batch(chunk)
            ops.raise_status(device=self.X.device)
            Y = model.detect_axons(Xb.flatten(0, 1))
            heads.append(Y.reshape(len(chunk), self.ntiles, *Y.shape[1:]))
        return torch.cat(heads) if heads else torch.empty((0, self.ntiles, 0, 0, 3), device=self.X.device)
