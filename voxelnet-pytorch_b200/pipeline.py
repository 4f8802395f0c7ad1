"""The fused front-end fast path: raw points -> (B,128,H,W) RPN input, everything on the device.

This is the sequence predict.py:52-60 / RPN3D.forward (model.py:298-306) run in the reference -
pcl_to_voxels -> prepare_voxel -> H2D -> feature_net -> middle_layer -> reshape - without the
host round trips: raw points are voxelized on the GPU, the voxel count never leaves the
device, the dense grid is a resident bf16 buffer that is written and re-zeroed only at the
occupied rows, and the three Conv3d layers chain in bf16 NDHWC.

All buffers are allocated once per (batch, capacity) shape; `forward_points` only enqueues
kernels on the current stream (no host sync, CUDA-graph capturable).
"""
from __future__ import annotations

import torch

from . import _cabi, ops
from .config import grid_config
from .model import FeatureLearningNet, MiddleConvNet
from .utils import VoxelBatch, pack_frames


class StageTimer:
    """CUDA-event stopwatch around the stages of `forward_points` (events are recorded on the
    launching stream; reading them is deferred to `summary()`, after a synchronize)."""

    def __init__(self):
        self.records = []  # (stage, start_event, end_event)

    def span(self, stage: str):
        return _Span(self, stage)

    def summary(self) -> dict:
        """stage -> (mean ms per recorded span, spans)."""
        acc = {}
        for stage, e0, e1 in self.records:
            t, n = acc.get(stage, (0.0, 0))
            acc[stage] = (t + e0.elapsed_time(e1), n + 1)
        return {k: (t / n, n) for k, (t, n) in acc.items()}


class _Span:
    def __init__(self, timer, stage):
        self.timer, self.stage = timer, stage

    def __enter__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.e0.record()

    def __exit__(self, *exc):
        self.e1.record()
        self.timer.records.append((self.stage, self.e0, self.e1))


class _NoSpan:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_SPAN = _NoSpan()


class VoxelNetFrontEnd:
    def __init__(self, feature_net: FeatureLearningNet, middle_net: MiddleConvNet, target: str = "Car",
                 voxels_per_frame_capacity: int = 49152, fuse_voxel_features: bool = True,
                 skip_empty_boxes: bool = True, resident_outputs: bool = True, materialise_grid=None,
                 grid_budget_bytes: int = 16 << 30, output_format: str = "nchw_f32"):
        self.feature_net, self.middle_net = feature_net, middle_net
        self.target = target
        self.cfg = grid_config(target)
        self.grid = _cabi.grid_for_target(target)
        self.cap_per_frame = int(voxels_per_frame_capacity)
        # True: the (K,T,7) feature buffer is never materialised (vn_voxelize_svfe_fused);
        # False: voxel dict written then read back by the SVFE kernel (the reference's data flow)
        self.fuse_voxel_features = bool(fuse_voxel_features)
        # The sparse -> dense step (model.py:102-106) of the eval fast path has two bit-identical forms:
        #   False: the (B,D,H,W,128) tensor is NEVER materialised.  The SVFE kernel writes a compact (K,128) bf16 row
        #          table and the first conv layer builds its operand boxes from it through the voxelizer's occupancy
        #          index (vn_voxelize_svfe_rows / vn_conv3d_bn_relu_fwd_rows): 0.2 GB instead of 5.8 GB at Car batch 16,
        #          0.30 instead of 1.04 GB of DRAM traffic in conv-1, no scatter / re-zero pass - but conv-1 itself is
        #          ~25 % slower (boxes built by warps instead of one TMA each).
        #   True:  a resident, zero-between-steps bf16 grid that is written at the K occupied cells, read by TMA and
        #          re-zeroed (the round-1 form; ~3 % more frames/s at Car batch 16).
        #   None (default): the grid if it fits `grid_budget_bytes` (16 GiB: Car up to 44 frames per batch), else rows.
        # The voxel-dict data flow (fuse_voxel_features=False) and the training step always use the grid.
        # "nchw_f32": the reference's (B,128,H,W) float32 tensor, channel c*2+d (model.py:262) - the drop-in.
        # "nhwc_bf16": the same values rounded to bf16 in channels-last memory, returned as a (B,128,H,W) view whose
        #   channel k = d*64 + c holds the reference's channel c*2 + d (`reference_channel_of_output()`): for an RPN that
        #   runs in bf16 channels-last - permute the input channels of its first convolution once
        #   (`weight[:, front_end.reference_channel_of_output()]`) and the two feed it bit for bit the same numbers, with
        #   half the bytes written by the last conv layer and no cast / layout pass in between.
        if output_format not in ("nchw_f32", "nhwc_bf16"):
            raise ValueError("output_format must be 'nchw_f32' or 'nhwc_bf16'")
        self.output_format = output_format
        self._grid_request = None if materialise_grid is None else bool(materialise_grid)
        self.grid_budget_bytes = int(grid_budget_bytes)
        self.materialise_grid = True if not self.fuse_voxel_features else self._grid_request  # None until the batch size is known
        # True: the middle layers skip input boxes that hold only background.  conv-1's input is the
        # ~99 % empty voxel grid (exact skip); conv-2/3 then read their inputs as deviations from the
        # constant the previous layer outputs over empty space, which are zero wherever no voxel is in
        # reach (vn_conv3d_fold_background / vn_conv3d_mark_boxes).  False: dense, plain activations.
        self.skip_empty_boxes = bool(skip_empty_boxes)
        # True (needs skip_empty_boxes): the two intermediate activations and the returned tensor are
        # resident buffers with a per-block state table; a step stores only the 16x8 output blocks a
        # voxel can reach now or could reach in the previous step (those are rewritten with the layer's
        # background constant).  Everything else already holds that constant - bit-identical result.
        self.resident_outputs = bool(resident_outputs) and self.skip_empty_boxes
        # True: conv-2/3 planning and the grid re-zeroing run on a side stream next to conv-1 / conv-2 (when no
        # StageTimer is attached: stage times are measured on the serial launch order)
        self.overlap_small_kernels = True
        self._shape = None
        # bumped whenever the device buffers are (re)allocated: CUDA graphs captured before hold stale pointers
        self._generation = 0
        self._layer_geometry = [(l.stride[0], l.padding[0]) for l in middle_net.middle_layer]
        self.refresh_parameters()

    def _param_signature(self) -> tuple:
        """(storage, version) of every parameter and buffer the folded tables are derived from: in-place updates
        (optimizer steps, load_state_dict) bump `_version`, re-assignment changes the pointer."""
        sig = []
        for mod in (self.feature_net, self.middle_net.middle_layer):
            for t in list(mod.parameters()) + list(mod.buffers()):
                sig.append((t.data_ptr(), t._version))
        return tuple(sig)

    def refresh_parameters(self) -> None:
        """Re-fold BN / re-pack weights after the modules' parameters changed.  `forward_points` calls this by itself
        when a parameter or BatchNorm buffer was modified since the last fold (checked by tensor version).  The folded
        tables are new tensors, so steps captured before (`capture_graph`) become invalid: the generation is bumped
        and their `replay()` raises instead of reading freed tables."""
        self._sig = self._param_signature()
        self._generation += 1
        if self.feature_net.training or self.middle_net.training:
            # training mode folds nothing ahead of time: BatchNorm uses the statistics of each batch
            # (forward_points_train); the eval-mode fast path is set up by the next call after .eval()
            self._p1 = None
            return
        self._p1 = self.feature_net.vfe_1.folded()
        self._p2 = self.feature_net.vfe_2.folded()
        self._conv = [layer.folded() for layer in self.middle_net.middle_layer]
        # background chain: shift tables with the background's contribution folded in, and each
        # layer's own output background (subtracted on store, so the stored deviation is sparse)
        self._tables, self._bg = [], []
        if self.skip_empty_boxes:
            # layer 1 reads the zero grid: its output background relu(shift) is the same at every position.  Layer 2's own
            # background depends on the border class of the position (the reference zero-pads), and so does what it subtracts
            # on store: the stored deviation is exactly zero wherever no voxel is in reach, borders included, and layer 3 folds
            # the per-class background of its input into a 5 x 5-variant shift table - no layer has to treat the H/W border
            # ring as live.
            layers = list(self.middle_net.middle_layer)
            din = self.cfg.depth
            t1, bg1 = ops.conv3d_fold_background(layers[0].conv.weight, din, layers[0].stride[0], layers[0].padding[0],
                                                 self._conv[0][1], self._conv[0][2], None)
            bg1 = bg1[0, 0].contiguous()                      # (D1, 64)
            din = bg1.shape[0]
            t2, bg2 = ops.conv3d_fold_background(layers[1].conv.weight, din, layers[1].stride[0], layers[1].padding[0],
                                                 self._conv[1][1], self._conv[1][2], bg1)      # (3,3,D2,64) each
            din = bg2.shape[2]
            t3, _ = ops.conv3d_fold_background(layers[2].conv.weight, din, layers[2].stride[0], layers[2].padding[0],
                                               self._conv[2][1], self._conv[2][2], bg2)        # (5,5,D3,64)
            self._tables, self._bg = [t1, t2, t3], [bg1, bg2]
        if getattr(self, "out_states", None) is not None:
            for st in self.out_states:  # the constants the untouched blocks hold depend on the parameters
                st.fill_(0xFF)

    def reserve(self, n_frames: int, max_points: int, device=None) -> None:
        """Size every device buffer once for batches of `n_frames` frames and up to `max_points` points in total.
        Later batches that fit reuse the buffers (and the CUDA graphs captured on them); a larger batch reallocates
        everything, bumps `generation` and thereby invalidates every step captured before."""
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._ensure_buffers(int(n_frames), int(max_points), dev)

    @property
    def generation(self) -> int:
        return self._generation

    def _ensure_buffers(self, n_frames: int, n_points: int, device) -> None:
        shape = (n_frames, n_points, str(device))
        if self._shape is not None and self._shape[0] == n_frames and self._shape[1] >= n_points and self._shape[2] == str(device):
            return
        if self._shape is not None:
            # drop the old set first: two ~10 GB buffer sets per front end must not coexist
            for name in ("workspace", "voxels", "voxelwise", "dense", "rows", "_out_store", "box_flags", "conv_workspaces", "out_states", "act1", "act2", "out"):
                setattr(self, name, None)
            self._train_coords = None
        self._generation += 1
        self._pending_frames = n_frames
        c = self.cfg
        cap = min(self.cap_per_frame * n_frames, max(n_points, 1))
        T = c.points_per_voxel
        B, D, H, W = n_frames, c.depth, c.height, c.width
        d1 = ops.conv3d_out_depth(D, 2, 1)
        d2 = ops.conv3d_out_depth(d1, 1, 0)
        d3 = ops.conv3d_out_depth(d2, 2, 1)
        self.workspace = torch.empty(ops.voxelize_workspace_bytes(self.grid, n_frames, n_points), dtype=torch.uint8, device=device)
        self.voxels = VoxelBatch(
            torch.empty((0 if self.fuse_voxel_features else cap, T, 7), dtype=torch.float32, device=device),
            torch.empty((cap, 4), dtype=torch.int64, device=device),
            torch.empty((cap,), dtype=torch.int64, device=device),
            torch.zeros(n_frames + 1, dtype=torch.int32, device=device),
            n_frames,
        )
        self.voxelwise = torch.empty((0 if self.fuse_voxel_features else cap, 128), dtype=torch.float32, device=device)
        self.dense = None
        self.rows = None
        if not self.fuse_voxel_features:
            self.materialise_grid = True
        elif self._grid_request is None:
            self.materialise_grid = B * D * H * W * 128 * 2 <= self.grid_budget_bytes
        if self.materialise_grid:
            self._ensure_dense(device)
        else:
            self.rows = torch.empty((cap, 128), dtype=torch.bfloat16, device=device)
        # live-box tables of the three layers: set by mark_boxes from the voxel coordinates, reset by
        # the layer that consumes them
        self.box_flags = [ops.conv3d_new_box_flags(B, d, H, W, device) for d in (D, d1, d2)] if self.skip_empty_boxes else None
        # one plan per layer (kept after the step so that diagnostics can decode what was executed)
        self.conv_workspaces = [torch.empty(ops.conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=device) for _ in range(3)]
        self.out_states = [ops.conv3d_new_out_state(B, H, W, device) for _ in range(3)] if self.resident_outputs else None
        self.act1 = torch.empty((B, d1, H, W, 64), dtype=torch.bfloat16, device=device)
        self.act2 = torch.empty((B, d2, H, W, 64), dtype=torch.bfloat16, device=device)
        if self.output_format == "nhwc_bf16":
            self._out_store = torch.empty((B, H, W, 64 * d3), dtype=torch.bfloat16, device=device)
            self.out = self._out_store.permute(0, 3, 1, 2)  # (B,128,H,W), channels-last strides
        else:
            self._out_store = self.out = torch.empty((B, 64 * d3, H, W), dtype=torch.float32, device=device)
        self.capacity = cap
        self._shape = shape

    def _ensure_dense(self, device) -> None:
        """The resident bf16 grid (materialise_grid, the voxel-dict data flow, the training step): zero-filled once,
        every step re-zeroes the cells it wrote."""
        if getattr(self, "dense", None) is None:
            c = self.cfg
            self.dense = torch.empty((self._pending_frames, c.depth, c.height, c.width, 128), dtype=torch.bfloat16, device=device)
            ops.fill_zero(self.dense)

    @property
    def _final(self):
        return "nhwc_bf16" if self.output_format == "nhwc_bf16" else True

    def reference_channel_of_output(self) -> torch.Tensor:
        """For every channel k of the returned tensor the channel of the reference's (B,128,H,W) tensor it holds
        (identity for "nchw_f32"; for "nhwc_bf16" k = d*64 + c holds reference channel c*D3 + d)."""
        d3 = ops.conv3d_out_depth(ops.conv3d_out_depth(ops.conv3d_out_depth(self.cfg.depth, 2, 1), 1, 0), 2, 1)
        k = torch.arange(64 * d3)
        return k if self.output_format == "nchw_f32" else (k % 64) * d3 + k // 64

    def device_bytes(self) -> dict:
        """Bytes of device memory this front end holds, by buffer (after the first call / reserve())."""
        out = {}
        for name in ("workspace", "dense", "rows", "act1", "act2", "out"):
            t = getattr(self, name, None)
            out[name] = 0 if t is None else t.numel() * t.element_size()
        out["coordinates"] = self.voxels.coordinate.numel() * 8 if getattr(self, "voxels", None) is not None else 0
        out["plans_flags_states"] = sum(t.numel() * t.element_size() for lst in (self.box_flags, self.conv_workspaces, self.out_states)
                                        if lst is not None for t in lst)
        out["total"] = sum(out.values())
        return out

    def forward_points(self, points: torch.Tensor, frame_offsets: torch.Tensor, max_points_per_frame: int,
                       timer: StageTimer = None) -> torch.Tensor:
        """points (N,4) f32 cuda (frames concatenated), frame_offsets (B+1,) i32 cuda ->
        (B,128,H,W) f32.  The returned tensor is an internal buffer reused by the next call; with
        resident_outputs it must not be modified by the caller (blocks no voxel reaches are not rewritten)."""
        if self.feature_net.training or self.middle_net.training:
            raise RuntimeError("forward_points is the eval-mode fast path; use forward_points_train in training mode")
        if self._p1 is None or self._sig != self._param_signature():
            self.refresh_parameters()
        n_frames = frame_offsets.numel() - 1
        self._ensure_buffers(n_frames, points.shape[0], points.device)
        self._clear_train_cells()
        vb = self.voxels
        span = timer.span if timer is not None else (lambda stage: _NO_SPAN)
        k_dev = vb.n_voxels_dev
        overlap = self.skip_empty_boxes and timer is None and self.overlap_small_kernels
        early = overlap and self.fuse_voxel_features
        use_rows = not self.materialise_grid
        if not use_rows:
            self._ensure_dense(points.device)
        n_pts = points.shape[0]
        if early:
            # coordinates right after the bucketing; live-box marking and the three conv plans then run on the side
            # stream next to the SVFE kernel
            ops.voxelize_index(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, self.workspace)
            ops.voxelize_bucket_coords(self.grid, points, frame_offsets, vb.voxel_offsets, vb.coordinate, self.workspace)
            main = torch.cuda.current_stream(points.device)
            aux = self._aux_stream(points.device)
            ev = self._events
            ev[4].record(main)
            with torch.cuda.stream(aux):
                aux.wait_event(ev[4])
                ops.conv3d_mark_boxes(vb.coordinate, n_frames, self.cfg.depth, self.cfg.height, self.cfg.width,
                                      self._layer_geometry, self.box_flags, n_voxels_dev=k_dev)
                st_ = self.out_states if self.out_states is not None else (None, None, None)
                B_, H_, W_ = n_frames, self.cfg.height, self.cfg.width
                ops.conv3d_plan(B_, self.cfg.depth, H_, W_, 2, 1, self.box_flags[0], self.conv_workspaces[0], consume_flags=True,
                                out_state=st_[0])
                ops.conv3d_plan(B_, self.act1.shape[1], H_, W_, 1, 0, self.box_flags[1], self.conv_workspaces[1],
                                consume_flags=True, out_state=st_[1])
                ops.conv3d_plan(B_, self.act2.shape[1], H_, W_, 2, 1, self.box_flags[2], self.conv_workspaces[2],
                                consume_flags=True, out_state=st_[2])
                ev[1].record(aux)
            if use_rows:
                ops.voxelize_svfe_rows(self.grid, points, frame_offsets, vb.voxel_offsets, self._p1, self._p2, self.rows,
                                       vb.coordinate, self.workspace, bucketed=True)
            else:
                ops.voxelize_svfe_fused(self.grid, points, frame_offsets, vb.voxel_offsets, self._p1, self._p2,
                                        None, vb.coordinate, None, self.workspace, dense=self.dense, bucketed=True)
        elif self.fuse_voxel_features:
            with span("voxelize"):  # crop, cell keys, occupancy bitmap, ranks, per-voxel buckets
                ops.voxelize_index(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, self.workspace)
            with span("svfe"):      # bucket -> first-T selection -> 7-d rows on chip -> VFE-1/VFE-2/max -> row table / grid cell
                if use_rows:
                    ops.voxelize_svfe_rows(self.grid, points, frame_offsets, vb.voxel_offsets, self._p1, self._p2, self.rows,
                                           vb.coordinate, self.workspace)
                else:
                    ops.voxelize_svfe_fused(self.grid, points, frame_offsets, vb.voxel_offsets, self._p1, self._p2,
                                            None, vb.coordinate, None, self.workspace, dense=self.dense)
        else:
            with span("voxelize"):
                ops.voxelize_index(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, self.workspace)
                ops.voxelize_fill(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, vb.feature,
                                  vb.coordinate, vb.number, self.workspace)
            with span("svfe"):
                ops.svfe_fused_fwd(vb.feature, self._p1, self._p2, out=self.voxelwise, n_voxels_dev=k_dev)
        with span("scatter"):
            if not self.fuse_voxel_features:  # (the fused kernel wrote the grid cells itself)
                ops.scatter_dense_bf16(self.voxelwise, vb.coordinate, self.dense, clear=False, n_voxels_dev=k_dev)
            if self.skip_empty_boxes and not early:
                ops.conv3d_mark_boxes(vb.coordinate, n_frames, self.cfg.depth, self.cfg.height, self.cfg.width,
                                      self._layer_geometry, self.box_flags, n_voxels_dev=k_dev)
        (w0, s0, t0), (w1, s1, t1), (w2, s2, t2) = self._conv
        ws0, ws1, ws2 = self.conv_workspaces
        st0, st1, st2 = self.out_states if self.out_states is not None else (None, None, None)

        def conv1(**kw):  # the first layer: from the row table (no grid) or from the resident grid
            if use_rows:
                return ops.conv3d_bn_relu_fwd_rows(self.grid, self.rows, vb.voxel_offsets, n_pts, self.workspace, w0, s0, t0, 2, 1,
                                                   out=self.act1, **kw)
            return ops.conv3d_bn_relu_fwd(self.dense, w0, s0, t0, 2, 1, out=self.act1, **kw)

        def clear_grid():  # re-zero the grid cells of this step (nothing to do without a grid)
            if not use_rows:
                ops.scatter_dense_bf16(None, vb.coordinate, self.dense, clear=True, n_voxels_dev=k_dev)

        if overlap:
            # Same launches as the serial branch below, with the small latency-bound kernels off the critical path:
            # the work plans of conv-2/3 depend on the voxel coordinates only and are built on a side stream while
            # conv-1 runs; the re-zeroing of the grid cells runs there next to conv-2.  (Fork / join by events: a
            # CUDA graph captured from this call keeps the branches parallel.)
            f0, f1, f2 = self.box_flags
            B, H, W = n_frames, self.cfg.height, self.cfg.width
            main = torch.cuda.current_stream(points.device)
            aux = self._aux_stream(points.device)
            ev = self._events
            if early:  # all three plans were built next to the SVFE kernel
                main.wait_event(ev[1])
            else:
                ev[0].record(main)
                aux.wait_event(ev[0])
                with torch.cuda.stream(aux):
                    ops.conv3d_plan(B, self.act1.shape[1], H, W, 1, 0, f1, ws1, consume_flags=True, out_state=st1)
                    ops.conv3d_plan(B, self.act2.shape[1], H, W, 2, 1, f2, ws2, consume_flags=True, out_state=st2)
                    ev[1].record(aux)
            conv1(box_flags=f0, consume_flags=True, sub=self._bg[0], workspace=ws0, out_state=st0, planned=early)
            if not use_rows:
                ev[2].record(main)
                with torch.cuda.stream(aux):
                    aux.wait_event(ev[2])
                    clear_grid()
                    ev[3].record(aux)
            if not early:
                main.wait_event(ev[1])
            ops.conv3d_bn_relu_fwd(self.act1, w1, s1, self._tables[1], 1, 0, out=self.act2, box_flags=f1, reverse=True,
                                   consume_flags=True, sub=self._bg[1], workspace=ws1, out_state=st1, planned=True)
            ops.conv3d_bn_relu_fwd(self.act2, w2, s2, self._tables[2], 2, 1, final=self._final, out=self._out_store, box_flags=f2,
                                   consume_flags=True, workspace=ws2, out_state=st2, planned=True)
            if not use_rows:
                main.wait_event(ev[3])  # join: the next step's grid writes come after the clear
        elif self.skip_empty_boxes:
            f0, f1, f2 = self.box_flags
            B_, H_, W_ = n_frames, self.cfg.height, self.cfg.width
            with span("plan"):  # the three device-built work plans (in the overlapped order they run on the side stream)
                ops.conv3d_plan(B_, self.cfg.depth, H_, W_, 2, 1, f0, ws0, consume_flags=True, out_state=st0)
                ops.conv3d_plan(B_, self.act1.shape[1], H_, W_, 1, 0, f1, ws1, consume_flags=True, out_state=st1)
                ops.conv3d_plan(B_, self.act2.shape[1], H_, W_, 2, 1, f2, ws2, consume_flags=True, out_state=st2)
            with span("conv1"):  # zero background in, deviation from relu(shift) out
                conv1(box_flags=f0, consume_flags=True, sub=self._bg[0], workspace=ws0, out_state=st0, planned=True)
            with span("unscatter"):
                clear_grid()
            with span("conv2"):
                ops.conv3d_bn_relu_fwd(self.act1, w1, s1, self._tables[1], 1, 0, out=self.act2, box_flags=f1, reverse=True,
                                       consume_flags=True, sub=self._bg[1], workspace=ws1, out_state=st1, planned=True)
            with span("conv3"):
                ops.conv3d_bn_relu_fwd(self.act2, w2, s2, self._tables[2], 2, 1, final=self._final, out=self._out_store, box_flags=f2,
                                       consume_flags=True, workspace=ws2, out_state=st2, planned=True)
        else:
            with span("conv1"):
                conv1(workspace=ws0)
            with span("unscatter"):
                clear_grid()
            with span("conv2"):
                ops.conv3d_bn_relu_fwd(self.act1, w1, s1, t1, 1, 0, out=self.act2, workspace=ws1)
            with span("conv3"):
                ops.conv3d_bn_relu_fwd(self.act2, w2, s2, t2, 2, 1, final=self._final, out=self._out_store, workspace=ws2)
        return self.out

    def _aux_stream(self, device):
        if getattr(self, "_aux", None) is None or self._aux.device != device:
            self._aux = torch.cuda.Stream(device=device)
            self._events = [torch.cuda.Event() for _ in range(5)]
        return self._aux

    def capture_graph(self, points: torch.Tensor, frame_offsets: torch.Tensor, max_points_per_frame: int):
        """Capture one eval-mode step on these (static) input buffers into a CUDA graph; `graph.replay()` then
        re-runs the whole voxelize -> SVFE -> grid -> 3 x conv sequence with one launch and writes `self.out`.
        The step is host-sync free and its bookkeeping (voxel counts, live-box tables, block state) lives on the
        device, so replays of graphs captured for different input buffers can be interleaved freely."""
        for _ in range(2):  # warm-up outside the capture: buffers, TMA descriptors, function attributes
            self.forward_points(points, frame_offsets, max_points_per_frame)
        torch.cuda.current_stream().synchronize()
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=points.device)
        with torch.cuda.graph(graph, stream=side):
            self.forward_points(points, frame_offsets, max_points_per_frame)
        return CapturedStep(self, graph)

    def _clear_train_cells(self) -> None:
        """Re-zero the grid cells the previous training step wrote (the eval path clears its own)."""
        prev = getattr(self, "_train_coords", None)
        if prev is not None and prev.shape[0] > 0 and self.dense is not None:
            ops.scatter_dense_bf16(None, prev, self.dense, clear=True)
        self._train_coords = None

    def forward_points_train(self, points: torch.Tensor, frame_offsets: torch.Tensor, max_points_per_frame: int,
                             timer: StageTimer = None) -> torch.Tensor:
        """Training-mode step (model.train(True), train.py:148): the same path with batch-statistics BatchNorm
        and an autograd graph.  points (N,4) f32 cuda, frame_offsets (B+1,) i32 cuda -> (B,128,H,W) f32 with a
        grad_fn; `.backward()` on anything computed from it fills the .grad of every VFE / ConvMD parameter.

        Differences from the eval path: the (K,T,7) feature buffer is materialised (the batch statistics need
        two reduction passes over all rows before the VFE can be applied, and the backward re-reads it), the
        exact voxel count is read back once per step, and every layer keeps its raw convolution output.  The
        resident grid keeps this step's voxel rows until the next call (the weight gradient of the first conv
        layer reads them in the backward), so call backward before the next forward."""
        if not (self.feature_net.training and self.middle_net.training):
            raise RuntimeError("forward_points_train needs both modules in .train() mode")
        n_frames = frame_offsets.numel() - 1
        self._ensure_buffers(n_frames, points.shape[0], points.device)
        self._ensure_dense(points.device)  # the training kernels (sparse wgrad of layer 1) read the resident grid
        self._clear_train_cells()
        self._p1 = None  # running statistics (and soon the weights) change: the eval path re-folds on its next call
        if self.out_states is not None:
            for st in self.out_states:  # the eval path's resident outputs are not maintained here
                st.fill_(0xFF)
        c, vb = self.cfg, self.voxels
        span = timer.span if timer is not None else (lambda stage: _NO_SPAN)
        with span("voxelize"):
            ops.voxelize_index(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, self.workspace)
            K = self.voxel_counts()[-1]  # blocking: the statistics are over exactly K*T rows
            if K == 0:  # nn.BatchNorm1d raises on an empty batch as well (model.py:76 in train mode)
                raise RuntimeError("forward_points_train: the batch has no voxel inside the crop range")
            feat = getattr(self, "_train_feature", None)
            if feat is None or feat.shape[0] < K or feat.device != points.device:
                feat = torch.empty((max(K, 1) * 5 // 4, c.points_per_voxel, 7), dtype=torch.float32, device=points.device)
                self._train_feature = feat
            ops.voxelize_fill(self.grid, points, frame_offsets, max_points_per_frame, vb.voxel_offsets, feat,
                              vb.coordinate, vb.number, self.workspace)
        feature, coords = feat[:K], vb.coordinate[:K].clone()
        with span("svfe"):
            vw = self.feature_net.voxelwise(feature)
        with span("scatter"):
            ops.scatter_dense_bf16(vw.detach(), coords, self.dense)
            self._train_coords = coords
        dense = self.dense.view(self.dense.shape)  # a fresh tensor object for the side channel
        if vw.requires_grad:
            dense._vn_sparse = (vw, coords)
        with span("conv"):
            return self.middle_net(dense)

    def plan_stats(self) -> list:
        """Blocking diagnostics: what the three conv launches of the last step executed (live boxes,
        tcgen05 FLOPs, blocks stored) - see ops.conv3d_plan_stats."""
        c = self.cfg
        B = self._shape[0]
        out, din = [], c.depth
        for layer, ws in zip(self.middle_net.middle_layer, self.conv_workspaces):
            out.append(ops.conv3d_plan_stats(ws, B, din, c.height, c.width, layer.cin, layer.stride[0], layer.padding[0]))
            din = ops.conv3d_out_depth(din, layer.stride[0], layer.padding[0])
        return out

    def forward_frames(self, frames: list) -> torch.Tensor:
        """Host frames [(N_i,4) f32 numpy] -> (B,128,H,W) f32 (includes the H2D copy)."""
        points, offsets, max_n = pack_frames(frames)
        return self.forward_points(points, offsets, max_n)

    def voxel_counts(self) -> list:
        """Blocking: per-frame voxel prefix of the last call; raises if the capacity overflowed."""
        offs = ops.voxel_offsets_to_host(self.voxels.voxel_offsets)
        if offs[-1] > self.capacity:
            raise RuntimeError("voxel capacity %d exceeded (%d voxels): raise voxels_per_frame_capacity" % (self.capacity, offs[-1]))
        return offs


class CapturedStep:
    """A CUDA graph of one eval-mode step, tied to the buffer generation of its front end: `replay()` refuses to run
    once the front end reallocated its buffers (a batch larger than any seen before, see `VoxelNetFrontEnd.reserve`)
    - the captured kernels would write through stale pointers."""

    def __init__(self, front_end, graph):
        self.front_end, self.graph, self.generation = front_end, graph, front_end.generation

    @property
    def valid(self) -> bool:
        return self.generation == self.front_end.generation

    def replay(self) -> None:
        if not self.valid:
            raise RuntimeError("captured step is stale: the front end reallocated its buffers (generation %d -> %d); "
                               "call reserve() with the largest batch first and capture again" % (self.generation, self.front_end.generation))
        self.graph.replay()


class HostFrameStream:
    """End-to-end feeding of a VoxelNetFrontEnd from pinned host batches (`dataset.PinnedFrameBatch`).

    Two device staging slots: `upload(batch)` enqueues the host->device copy of a batch on a copy stream,
    `forward()` runs the front end on the oldest uploaded batch on the current stream, so the copy of
    batch i+1 overlaps the compute of batch i.  Replaces the reference's per-element `.to(device)` of
    voxelized features (model.py:302-303) - only raw points cross PCIe.

        stream = HostFrameStream(front_end, n_frames=16, max_points=2_100_000)
        stream.upload(b0)
        for nxt in batches[1:] + [None]:
            if nxt is not None: stream.upload(nxt)
            out = stream.forward()                # (B,128,H,W) on the device, for the RPN
            counts = stream.voxel_counts_host()   # blocks until this step is done
    """

    def __init__(self, front_end, n_frames: int, max_points: int, device=None, cuda_graphs: bool = False):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        # front_end: one VoxelNetFrontEnd, or a list of P of them (sharing the same modules) to keep P steps in
        # flight: step s runs on front end s % P on that pipeline's own stream, so the short latency-bound kernels
        # of one step (index, bucketing) run beside the long convolution kernels of the other.  Each front end owns
        # its grid, workspaces and resident outputs.  forward() still hands every output to the CURRENT stream (the
        # current stream waits for the step), and as with one front end the returned buffer must be consumed - work
        # enqueued on the current stream - before the next forward() call.
        self.fes = list(front_end) if isinstance(front_end, (list, tuple)) else [front_end]
        self.fe, self.device = self.fes[0], dev
        for fe in self.fes:  # sized once for the largest batch the staging slots can hold: no reallocation under a graph
            fe.reserve(n_frames, max_points, dev)
        P = self.n_pipes = len(self.fes)
        self.compute = [torch.cuda.Stream(device=dev) for _ in range(P)] if P > 1 else [None]
        self._consumed = []  # events on the caller's stream, one per forward() call (the last P are kept)
        # cuda_graphs: an eval-mode step on a staging slot is captured into a CUDA graph the second time the same
        # (pipeline, slot, point count, max points per frame) comes by, and replayed from then on (fixed-size batches
        # in a steady stream hit this; anything else keeps running eagerly)
        self.cuda_graphs = bool(cuda_graphs)
        self._graphs, self._seen = {}, set()
        self.n_slots = P + 1  # one uploading while P compute
        self.slots = [(torch.empty((int(max_points), 4), dtype=torch.float32, device=dev),
                       torch.empty(n_frames + 1, dtype=torch.int32, device=dev)) for _ in range(self.n_slots)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.ready = [torch.cuda.Event() for _ in range(self.n_slots)]
        self.freed = [torch.cuda.Event() for _ in range(self.n_slots)]
        for e in self.freed:
            e.record(torch.cuda.current_stream(dev))
        self.pending = []  # (slot, n_points, max_points_per_frame)
        self.next_slot = 0
        # per-frame voxel prefix of the last P+1 steps (pinned), each with the event that marks its D2H copy done
        self._counts = [torch.zeros(n_frames + 1, dtype=torch.int32).pin_memory() for _ in range(P + 1)]
        self._counts_done = [torch.cuda.Event() for _ in range(P + 1)]
        self._steps = 0
        self.h2d_bytes = 0

    def upload(self, batch) -> None:
        if len(self.pending) >= self.n_slots:
            raise RuntimeError("every staging slot holds a batch that was not consumed: call forward() first")
        k = self.next_slot
        pts, offs = self.slots[k]
        batch.wait_filled()  # a pending fill_async must have landed in the pinned buffer
        if batch.n_points > pts.shape[0] or batch.offsets.numel() != offs.numel():
            raise ValueError("batch does not fit the staging slot (points %d > %d or frame count differs)" % (batch.n_points, pts.shape[0]))
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.freed[k])
            pts[: batch.n_points].copy_(batch.points[: batch.n_points], non_blocking=True)
            offs.copy_(batch.offsets, non_blocking=True)
            self.ready[k].record(self.copy_stream)
            # the host may overwrite the pinned batch once this copy is done (PinnedFrameBatch.fill waits for it)
            done = torch.cuda.Event()
            done.record(self.copy_stream)
            batch._in_flight = done
        self.h2d_bytes = batch.n_points * 16 + offs.numel() * 4
        self.pending.append((k, batch.n_points, batch.max_points_per_frame))
        self.next_slot = (k + 1) % self.n_slots

    def upload_frames(self, frames: list) -> None:
        """`upload` for a list of per-frame (N_i,4) float32 numpy arrays (what the reference's loader yields,
        dataset.py:51).  Arrays in page-locked memory (`dataset.pinned_frame`, `load_velodyne_bin(pinned=True)`) are
        copied straight into the device CSR slot, one asynchronous copy per frame and no host-side packing
        (vn_upload_frames); the arrays must then stay untouched until this batch's `forward()` has been called and a
        later `upload*` call reuses the slot (or the copy stream is synchronised).  If any array is ordinary pageable
        memory the batch goes through an internal ring of PinnedFrameBatch (host thread pool) instead."""
        import ctypes as C

        import numpy as np

        from . import _cabi

        if len(self.pending) >= self.n_slots:
            raise RuntimeError("every staging slot holds a batch that was not consumed: call forward() first")
        k = self.next_slot
        pts, offs = self.slots[k]
        n = offs.numel() - 1
        if len(frames) != n:
            raise ValueError("expected %d frames, got %d" % (n, len(frames)))
        fr = [f if (f.dtype == np.float32 and f.flags["C_CONTIGUOUS"]) else np.ascontiguousarray(f, dtype=np.float32) for f in frames]
        for i, f in enumerate(fr):
            if f.ndim != 2 or f.shape[1] != 4:
                raise ValueError("frame %d has shape %s, expected (N,4)" % (i, f.shape))
        if not all(torch.from_numpy(f).is_pinned() for f in fr if f.shape[0] > 0):
            if getattr(self, "_pack_ring", None) is None:
                from .dataset import PinnedFrameBatch

                self._pack_ring = [PinnedFrameBatch(n, pts.shape[0]) for _ in range(self.n_slots + 1)]
                self._pack_next = 0
            pb = self._pack_ring[self._pack_next]
            self._pack_next = (self._pack_next + 1) % len(self._pack_ring)
            return self.upload(pb.fill(fr))
        total = sum(int(f.shape[0]) for f in fr)
        if total > pts.shape[0]:
            raise ValueError("batch does not fit the staging slot (points %d > %d)" % (total, pts.shape[0]))
        if getattr(self, "_offs_host", None) is None:
            self._offs_host = [torch.zeros(n + 1, dtype=torch.int32).pin_memory() for _ in range(self.n_slots)]
            self._frames_alive = [None] * self.n_slots
        self.ready[k].synchronize()  # the previous copy out of this slot's pinned offsets is over
        ptrs = (C.c_void_p * n)(*[f.ctypes.data for f in fr])
        sizes = (C.c_int64 * n)(*[int(f.shape[0]) for f in fr])
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.freed[k])
            _cabi.check(_cabi.lib().vn_upload_frames(ptrs, sizes, n, pts.data_ptr(), pts.shape[0], self._offs_host[k].data_ptr(),
                                                     offs.data_ptr(), torch.cuda.current_stream().cuda_stream), "vn_upload_frames")
            self.ready[k].record(self.copy_stream)
        self._frames_alive[k] = fr
        self.h2d_bytes = total * 16 + offs.numel() * 4
        self.pending.append((k, total, max([int(f.shape[0]) for f in fr], default=0)))
        self.next_slot = (k + 1) % self.n_slots

    def forward(self, timer: StageTimer = None) -> torch.Tensor:
        if not self.pending:
            raise RuntimeError("forward() without an uploaded batch")
        k, n, max_n = self.pending.pop(0)
        pts, offs = self.slots[k]
        main = torch.cuda.current_stream(self.device)
        P = self.n_pipes
        pipe = self._steps % P
        fe = self.fes[pipe]
        cs = main
        if P > 1:
            cs = self.compute[pipe]
            # this step overwrites the buffers of step s-P (same front end); its output was consumed by work enqueued
            # before call s-P+1, i.e. before the event recorded at the start of that call
            ev = torch.cuda.Event()
            ev.record(main)
            self._consumed = (self._consumed + [ev])[-P:]
            if len(self._consumed) == P:
                cs.wait_event(self._consumed[0])
        cs.wait_event(self.ready[k])
        training = fe.feature_net.training
        if training and P > 1:
            raise RuntimeError("several front ends in flight are an eval-mode feature (they would race on the parameters' .grad)")
        key = (pipe, k, n, max_n, fe.generation)
        use_graph = self.cuda_graphs and not training and timer is None
        with torch.cuda.stream(cs):
            graph = self._graphs.get(key) if use_graph else None
            if graph is None and use_graph:
                for old in [q for q in self._graphs if q[0] == pipe and q[4] != fe.generation]:
                    del self._graphs[old]  # captured before the front end replaced buffers or tables
            if graph is None and use_graph and key in self._seen:
                graph = self._graphs[key] = fe.capture_graph(pts[:n], offs, max_n)
                if not graph.valid:  # (cannot happen after reserve(); kept as a guard)
                    raise RuntimeError("front end reallocated its buffers during graph capture")
            self._seen.add(key)
            if graph is not None:
                graph.replay()
                out = fe.out
            else:
                fn = fe.forward_points_train if training else fe.forward_points
                out = fn(pts[:n], offs, max_n, timer=timer)
            self.freed[k].record(cs)
            slot = self._steps % (P + 1)
            self._counts[slot].copy_(fe.voxels.voxel_offsets, non_blocking=True)
            self._counts_done[slot].record(cs)
        if P > 1:
            main.wait_event(self._counts_done[slot])  # the consumer's stream sees the finished step
        self._steps += 1
        return out

    @property
    def host_counts(self) -> torch.Tensor:
        return self._counts[(self._steps - 1) % (self.n_pipes + 1)]

    def voxel_counts_host(self, lag: int = 0) -> torch.Tensor:
        """The per-frame voxel prefix (B+1,) int32, in pinned host memory, of the last forward() (lag 0) or of an
        earlier one (lag up to the number of front ends; 1 with a single front end).  Blocks only until THAT step's device->host copy is done: with lag 1 the device queue
        never drains while the host reads results (one step of latency for the consumer)."""
        if not 0 <= lag <= self.n_pipes or self._steps - lag < 1:
            raise RuntimeError("voxel_counts_host(lag=%d): no such step (yet), or more than %d steps back" % (lag, self.n_pipes))
        slot = (self._steps - 1 - lag) % (self.n_pipes + 1)
        self._counts_done[slot].synchronize()
        counts = self._counts[slot]
        fe = self.fes[(self._steps - 1 - lag) % self.n_pipes]
        if int(counts[-1]) > fe.capacity:  # the kernels clamp: voxels beyond the capacity were dropped from the step
            raise RuntimeError("voxel capacity %d exceeded (%d voxels): raise voxels_per_frame_capacity" % (fe.capacity, int(counts[-1])))
        return counts
