#!/usr/bin/env python
"""Benchmark of the VoxelNet front-end hot path (BASELINE.json: frames/s voxelize -> VFE ->
Conv3d-middle, Car cfg).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config car16|ped16|stress64]

One "step" = one pass of voxelize -> SVFE -> scatter -> 3 x Conv3d(+BN+ReLU) over one batch of
synthetic KITTI-shaped frames (default: BASELINE.json configs[1], Car, batch 16 per GPU).  Inputs
are resident in HBM for `value`; `e2e` times the same step from pinned HOST point buffers through
the public API with the H2D copy and a D2H read of the per-frame voxel counts inside the timed
region.  Frames are independent, so N GPUs run N replicas on their own frames with no collective
(weak scaling); timing is CUDA events on the launching stream, max over ranks.

`--impl reference` times the reference's own CPU path on the host cores: the UNMODIFIED reference modules from
`baseline/_ref/` (or /root/reference) through oracle/ref_chain.py - `cpu_baseline.kind = "reference"` - and only if
that copy is absent the oracle port (oracle/cpu_chain.py, kind "port", reason stated).  Under torchrun only rank 0 runs
it.  Prints ONE JSON line on rank 0.

`--data kitti` replaces the synthetic frames by real ones from the reference's bundled KITTI drive
(baseline/_ref/data, placed there by oracle/install_ref.py) - the real-data cross-check.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNIT = "frames/s"


def metric_name(config: str) -> str:
    target = CONFIGS[config][0]
    return "frames/s voxelize->VFE->Conv3d-middle (%s cfg)" % ("Car" if target == "Car" else "Ped/Cyc")


CONFIGS = {
    # name: (target, batch per GPU, generator kwargs, description)
    "car16": ("Car", 16, {}, "VoxelNet Car forward, batch 16 synthetic KITTI-shaped frames (~120k pts) per GPU"),
    "ped16": ("Pedestrian", 16, {}, "VoxelNet Pedestrian/Cyclist forward, batch 16 per GPU"),
    "stress64": ("Car", 64, dict(azimuth_steps=4400, clutter=0.07), "dense-scene stress: ~300k pts/frame, batch 64 total"),
    # BASELINE.json configs[4]: hot-path training step (forward + backward + gradient all-reduce + SGD)
    "train16": ("Car", 16, {}, "VoxelNet Car training step of the hot path, batch 16 per GPU, NCCL gradient all-reduce"),
}

# dense-equivalent conv FLOPs per frame (padding taps counted)
_RESULT_FD = None


def quiet_stdout() -> None:
    """stdout carries exactly ONE line - the JSON result.  Libraries write there too (NCCL prints its version banner
    on stdout at communicator creation), so file descriptor 1 is pointed at stderr for the run and the result goes
    to a duplicate of the original stdout."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def conv_flops_per_frame(D, H, W):
    d1 = (D + 2 - 3) // 2 + 1
    d2 = d1 - 2
    d3 = (d2 + 2 - 3) // 2 + 1
    m = H * W
    return [2 * 27 * d1 * m * 128 * 64, 2 * 27 * d2 * m * 64 * 64, 2 * 27 * d3 * m * 64 * 64]


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=float(p["hbm_gbs"]), tflops=float(p["bf16_tflops_sustained"]), tflops_burst=float(p["bf16_tflops"]), source="measured")
    except Exception:
        # fallback stated in B200_PROFILING.md
        return dict(hbm_gbs=6650.0, tflops=1400.0, tflops_burst=1590.0, source="fallback")


class ClockSampler:
    """nvidia-smi clock / throttle sampler running during the timed region."""

    Q = "timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int, period_ms: int = 10):
        # 10 ms on a single GPU (a 50-step timed region is ~80 ms); 50 ms when several ranks share the host cores
        self.index, self.proc, self.lines, self.period_ms = index, None, [], int(period_ms)

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    @staticmethod
    def _epoch(ts: str):
        import datetime

        try:
            return datetime.datetime.strptime(ts.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
        except ValueError:
            return None

    def stop(self, t0: float = None, t1: float = None) -> dict:
        """Summary of the samples taken in the wall-clock window [t0, t1] (all samples if the
        window is empty or timestamps are unparseable)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 8:
                continue
            try:
                rows.append((self._epoch(parts[0]), float(parts[1]), float(parts[2]), float(parts[3]),
                             [n for n, v in zip(names, parts[4:8]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        inside = [r for r in rows if t0 is not None and r[0] is not None and t0 <= r[0] <= t1]
        use = inside if inside else rows
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        reasons = sorted({n for r in use for n in r[4]})
        return {"sm_mhz": statistics.median(r[1] for r in use), "sm_max_mhz": max(r[2] for r in use),
                "power_w_max": max(r[3] for r in use), "samples": len(use), "window": "timed region" if inside else "whole run",
                "reasons": reasons}


KITTI_DIRS = [os.path.join(ROOT, "baseline", "_ref", "data/2011_09_26/2011_09_26_drive_0001_sync/velodyne_points/data"),
              "/root/reference/data/2011_09_26/2011_09_26_drive_0001_sync/velodyne_points/data"]
DATA = "synthetic"


def kitti_frames(first: int, n: int) -> list:
    """n real frames of the reference's bundled KITTI drive starting at index `first` (wrapping over the files present)."""
    from voxelnet_pytorch_b200.dataset import load_velodyne_bin

    for d in KITTI_DIRS:
        if os.path.isdir(d):
            files = sorted(f for f in os.listdir(d) if f.endswith(".bin"))
            if files:
                return [load_velodyne_bin(os.path.join(d, files[(first + i) % len(files)])) for i in range(n)]
    raise SystemExit("bench.py --data kitti: no KITTI frames under baseline/_ref/data (run `python -m oracle.install_ref` in the container)")


def make_frames(cfg_name, rank, batch, seed=None):
    from voxelnet_pytorch_b200 import synth

    _, _, kw, _ = CONFIGS[cfg_name]
    seed0 = 1000 * rank if seed is None else seed
    if DATA == "kitti":
        return kitti_frames(seed0 % 997, batch)
    return synth.lidar_batch(batch, seed0=seed0, **kw)


def reference_chain():
    """(callable(frames, sd, target) -> (out, stage seconds), kind, note): the unmodified reference when a copy is
    reachable, else the oracle port."""
    from oracle import cpu_chain, ref_chain

    if ref_chain.available():
        from oracle import ref_import

        return (lambda fr, sd, tg: ref_chain.front_end_reference(fr, sd, tg, shuffle=True)), "reference", \
            "unmodified reference modules from %s (pcl_to_voxels with its shuffle -> prepare_voxel -> FeatureLearningNet -> middle_layer + reshape, predict.py:52-60)" % ref_import.REFERENCE_ROOT
    return (lambda fr, sd, tg: cpu_chain.front_end_cpu(fr, sd, tg)), "port", \
        "oracle port (baseline/_ref absent on this host: run `python -m oracle.install_ref` where /root/reference exists)"


def cpu_reference_sample(target, frames, sd, seconds_budget=20.0, max_frames=None):
    """Time the reference's CPU path (the reference itself, else the oracle port) one frame at a time within a time budget."""
    import torch

    chain, kind, note = reference_chain()
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))
    chain(frames[:1], sd, target)  # warm-up (thread pools, oneDNN primitives)
    times, stage_acc, n = [], {}, 0
    t_start = time.perf_counter()
    while True:
        f = frames[n % len(frames)]
        _, t = chain([f], sd, target)
        times.append(t["total"])
        for k, v in t.items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
        n += 1
        if (max_frames and n >= max_frames) or time.perf_counter() - t_start > seconds_budget or n >= 64:
            break
    best = min(times)
    return dict(frames=n, best_s=best, mean_s=sum(times) / n, stages_mean_s={k: v / n for k, v in stage_acc.items()},
                cores=torch.get_num_threads(), host_cpus=os.cpu_count(), kind=kind, note=note)


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path on the host cores, rank 0 only."""
    if rank != 0:
        return
    import torch
    from oracle import model_torch as OM

    # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host core it can get
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))

    target, batch, _, desc = CONFIGS[args.config]
    frames = make_frames(args.config, 0, min(batch, 4))
    sd = OM.random_state_dict(0)
    chain, kind, note = reference_chain()

    for _ in range(max(args.warmup, 1)):
        chain(frames[:1], sd, target)
    times = []
    for i in range(args.steps):
        _, t = chain([frames[i % len(frames)]], sd, target)
        times.append(t["total"])
    total = sum(times)
    value = args.steps / total  # each step is a bounded sample: ONE frame of the workload
    line = {
        "impl": "reference", "metric": metric_name(args.config), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": DATA,
        "config": {"workload": desc, "name": args.config, "sample": "1 frame per step (bounded sample of the batch)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                         "sample": "%d steps x 1 frame: %s" % (args.steps, note)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_train(args, rank, world, local_rank, dev):
    """Training step of the hot path (BASELINE.json configs[4]; reference loop train.py:146-155):
    model.train(True) forward with batch-statistics BatchNorm, backward from an upstream gradient of the
    (B,128,H,W) tensor (a fixed random tensor standing for the stock-PyTorch RPN + loss, which are out of scope),
    ONE flat all-reduce of the hot-path gradients over the ranks, clip_grad_norm_(5), SGD step, zero_grad.
    Points are resident in HBM; per-rank batch is fixed (weak scaling)."""
    import numpy as np
    import torch
    import torch.distributed as dist

    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200 import _cabi, sharding

    target, batch, _, desc = CONFIGS[args.config]
    torch.manual_seed(0)
    fnet, mnet = vn.FeatureLearningNet(target).to(dev).train(), vn.MiddleConvNet().to(dev).train()
    fe = vn.VoxelNetFrontEnd(fnet, mnet, target)
    params = list(fnet.parameters()) + list(mnet.parameters())
    opt = torch.optim.SGD(params, lr=1e-4)
    sets = []
    for seed in (None, 1000 * rank + 500):
        fr = make_frames("car16", rank, batch, seed=seed)
        sizes = [int(f.shape[0]) for f in fr]
        offs = np.zeros(batch + 1, dtype=np.int32)
        np.cumsum(sizes, out=offs[1:])
        sets.append((torch.from_numpy(np.concatenate(fr, 0)).to(dev), torch.from_numpy(offs).to(dev), max(sizes)))
    sets.sort(key=lambda x: -x[0].shape[0])
    c = fe.cfg
    dout = torch.randn((batch, 128, c.height, c.width), device=dev) * 1e-3
    spans = {k: [] for k in ("forward", "backward", "allreduce", "optimizer")}

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def step(i, record=False):
        pts, offs, max_n = sets[i % 2]
        t0 = ev()
        out = fe.forward_points_train(pts, offs, max_n)
        t1 = ev()
        out.backward(dout)
        t2 = ev()
        n = sharding.allreduce_gradients(params)
        t3 = ev()
        torch.nn.utils.clip_grad_norm_(params, 5.0)
        opt.step()
        opt.zero_grad(set_to_none=False)
        t4 = ev()
        if record:
            for k, a, b in (("forward", t0, t1), ("backward", t1, t2), ("allreduce", t2, t3), ("optimizer", t3, t4)):
                spans[k].append((a, b))
        return n

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    sampler = ClockSampler(local_rank, 10 if world == 1 else 50)
    sampler.start()
    step(0)
    barrier()
    wall0 = time.time()
    launches0 = _cabi.launch_count()
    e0 = ev()
    for i in range(args.steps):
        n_grad = step(i, record=True)
    e1 = ev()
    barrier()
    wall1 = time.time()
    launches = _cabi.launch_count() - launches0
    clocks = sampler.stop(wall0, wall1)
    ms = e0.elapsed_time(e1)
    stage = {k: sum(a.elapsed_time(b) for a, b in v) / max(len(v), 1) for k, v in spans.items()}
    ms_max = sharding.reduce_max([ms])[0]
    voxels = fe.voxel_counts()[-1]
    if rank != 0:
        return
    flops = sum(conv_flops_per_frame(c.depth, c.height, c.width))
    line = {
        "metric": "frames/s hot-path training step (Car cfg)", "value": world * batch * args.steps / (ms_max / 1e3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": DATA,
        "config": {"workload": desc, "name": args.config, "batch_per_gpu": batch, "voxels_per_batch": voxels,
                   "upstream_gradient": "fixed random (B,128,H,W) f32 tensor (RPN + loss are stock PyTorch, out of scope)",
                   "optimizer": "SGD + clip_grad_norm_(5) (train.py:146-155)", "grad_elements_allreduced": n_grad,
                   "l2_policy": "working set (dense grid, raw outputs, activations, gradients: > 15 GB) exceeds L2",
                   "vfe": "fp32", "conv": "bf16 operands, fp32 accumulate (tcgen05 forward, dgrad, wgrad)"},
        "stages_ms": stage,
        "dense_equivalent_conv_tflops": 3 * flops * batch / (ms_max / args.steps / 1e3) / 1e12,
        "clocks": clocks, "gpu_launches": launches,
    }
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="car16", choices=sorted(CONFIGS))
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget for the cpu_baseline leg (rank 0, N=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cuda-graph", action="store_true", help="time eager launches instead of CUDA-graph replays")
    ap.add_argument("--data", default="synthetic", choices=["synthetic", "kitti"],
                    help="kitti: real frames of the reference's bundled drive from baseline/_ref/data (cross-check)")
    ap.add_argument("--grid", default="auto", choices=["auto", "resident", "rows"],
                    help="sparse->dense form of the eval path: resident bf16 grid, never-materialised row table, or the library's default policy")
    ap.add_argument("--output", default="nchw_f32", choices=["nchw_f32", "nhwc_bf16"],
                    help="format of the returned tensor: the reference's float32 (B,128,H,W) (default), or bf16 channels-last for a bf16 RPN")
    ap.add_argument("--host-threads", type=int, default=None, help="threads of the host frame packer (default: cores / ranks, 2..8)")
    ap.add_argument("--no-rpn", action="store_true", help="skip the points -> RPN maps figure (detection_pipeline, N=1 only)")
    ap.add_argument("--in-flight", type=int, default=None,
                    help="eval steps in flight per GPU (front ends on their own streams; 1 = strictly one step at a time); "
                         "default 3, or 1 for batches above 16 frames per GPU (each pipeline holds its own grid and outputs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    global DATA
    DATA = args.data

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    quiet_stdout()

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL_DEBUG is the caller's; stdout is already pointed at stderr (quiet_stdout), so NCCL's banner cannot
        # reach the result line
        dist.init_process_group("nccl", device_id=dev)

    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200 import _cabi
    from voxelnet_pytorch_b200.pipeline import StageTimer

    target, batch_cfg, _, desc = CONFIGS[args.config]
    if args.config == "train16":
        run_train(args, rank, world, local_rank, dev)
        if world > 1:
            dist.destroy_process_group()
        return
    if args.config == "stress64":  # strong-scaling stress: 64 frames in total, frame i -> rank i mod world
        from voxelnet_pytorch_b200 import sharding as _sh

        mine = _sh.shard_frames(64, rank, world)
        batch = len(mine)
        frames = [make_frames(args.config, 0, 1, seed=1000 + i)[0] for i in mine]
    else:
        batch = batch_cfg
        frames = make_frames(args.config, rank, batch)

    torch.manual_seed(0)
    fnet = vn.FeatureLearningNet(target).eval()
    mnet = vn.MiddleConvNet().eval()
    # non-trivial BatchNorm statistics (random init would make BN the identity)
    g = torch.Generator().manual_seed(1)
    for m in list(fnet.modules()) + list(mnet.modules()):
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.weight.data = 0.5 + torch.rand(m.weight.shape, generator=g)
            m.bias.data = (torch.rand(m.bias.shape, generator=g) - 0.5) * 0.6
            m.running_mean.data = (torch.rand(m.running_mean.shape, generator=g) - 0.5) * 0.4
            m.running_var.data = 0.5 + torch.rand(m.running_var.shape, generator=g)
    grid_kw = dict({"auto": {}, "resident": {"materialise_grid": True}, "rows": {"materialise_grid": False}}[args.grid])
    if args.output != "nchw_f32":
        grid_kw["output_format"] = args.output
    fe = vn.VoxelNetFrontEnd(fnet.to(dev), mnet.to(dev), target, **grid_kw)

    # ---- inputs: pinned host staging + resident device copies of TWO different batches --------------
    # Steps alternate between them: the resident conv outputs are only rewritten where a voxel reaches
    # now or reached in the previous step, so repeating one batch would hide the restore traffic.
    import numpy as np

    if args.config == "stress64":
        frames_b = [make_frames(args.config, 0, 1, seed=5000 + i)[0] for i in mine]
    else:
        frames_b = make_frames(args.config, rank, batch, seed=1000 * rank + 500)

    class Batch:
        def __init__(self, fr):
            sizes = [int(f.shape[0]) for f in fr]
            self.offs_np = np.zeros(batch + 1, dtype=np.int32)
            np.cumsum(sizes, out=self.offs_np[1:])
            self.host_points = torch.from_numpy(np.concatenate(fr, 0)).pin_memory()
            self.host_offsets = torch.from_numpy(self.offs_np).pin_memory()
            self.max_n = max(sizes)
            self.points = self.host_points.to(dev)
            self.offsets = self.host_offsets.to(dev)

    batches = [Batch(frames), Batch(frames_b)]
    batches.sort(key=lambda x: -x.points.shape[0])  # the larger one first: buffers are sized once
    host_counts = torch.empty(batch + 1, dtype=torch.int32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i, timer=None):
        bt = batches[i % 2]
        return fe.forward_points(bt.points, bt.offsets, bt.max_n, timer=timer)

    # ---- device-resident timing ------------------------------------------------------------------
    counts = []
    for i in range(max(args.warmup, 2)):
        step(i)
        if i < 2:
            torch.cuda.synchronize()
            counts.append(fe.voxel_counts())  # also checks the voxel capacity
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank, 10 if world == 1 else 50)
    sampler.start()
    for i in range(2):  # keep the GPU busy while nvidia-smi spins up
        step(i)
    # One CUDA graph per alternating batch (the step is host-sync free); eager launches if capture fails.
    # --in-flight P > 1: P front ends (own grid, workspaces, resident outputs; shared modules), step i on front end
    # i % P on that pipeline's stream, so the latency-bound index / bucketing kernels of one step run beside the
    # convolution kernels of another.  Every front end keeps alternating between the two batches.
    P = max(1, args.in_flight) if args.in_flight is not None else (3 if batch <= 16 else 1)
    fes = [fe] + [vn.VoxelNetFrontEnd(fnet, mnet, target, **grid_kw) for _ in range(P - 1)]
    pipes = [torch.cuda.Stream(device=dev) for _ in range(P)] if P > 1 else [torch.cuda.current_stream(dev)]

    def batch_of(i):
        return (i // P + i % P) % 2

    graphs, launch_mode, per_step_launches = None, "eager launches", None
    if not args.no_cuda_graph:
        try:
            l0 = _cabi.launch_count()
            graphs = [[f.capture_graph(bt.points, bt.offsets, bt.max_n) for bt in batches] for f in fes]
            per_step_launches = (_cabi.launch_count() - l0) // (3 * len(batches) * P)  # 2 warm-ups + 1 captured run each
            launch_mode = "CUDA graph replay (one graph per front end and alternating batch)"
        except Exception as exc:  # noqa: BLE001 - fall back to the eager path, say so in the line
            graphs, launch_mode = None, "eager launches (graph capture failed: %s)" % type(exc).__name__
            torch.cuda.synchronize()

    def issue(i):
        with torch.cuda.stream(pipes[i % P]):
            if graphs is not None:
                graphs[i % P][batch_of(i)].replay()
            else:
                bt = batches[batch_of(i)]
                fes[i % P].forward_points(bt.points, bt.offsets, bt.max_n)

    for i in range(4 * P):
        issue(i)
    barrier()
    wall0 = time.time()
    launches0 = _cabi.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main_stream = torch.cuda.current_stream(dev)
    e0.record(main_stream)
    if P > 1:
        for ps in pipes:
            ps.wait_event(e0)
    for i in range(args.steps):
        issue(i)
    if P > 1:
        for ps in pipes:
            main_stream.wait_stream(ps)
    e1.record(main_stream)
    barrier()
    wall1 = time.time()
    launches = per_step_launches * args.steps if graphs is not None else _cabi.launch_count() - launches0
    clocks = sampler.stop(wall0, wall1)
    ms_total = e0.elapsed_time(e1)
    # per-stage times: a separate eager pass with CUDA events between the stages (the events cost ~2 % of a step,
    # so they stay out of the timed region above)
    timer = StageTimer()
    for i in range(min(args.steps, 20)):
        step(i, timer=timer)
    torch.cuda.synchronize()
    # what the conv launches executed (decoded from the device-built plans of the last two steps)
    plan = []
    for i in range(2):
        step(args.steps + i)
        torch.cuda.synchronize()
        plan.append(fe.plan_stats())
    plan = [{k: (plan[0][l][k] + plan[1][l][k]) / 2 for k in plan[0][l]} for l in range(3)]

    # ---- end to end: per-frame host arrays in, voxel counts out --------------------------------------
    # Starts where the reference starts: a list of per-frame (N_i, 4) float32 numpy arrays in ordinary (pageable)
    # host memory (voxelnet/dataset.py:51).  Inside the timed region, EVERY step: its 16 arrays are packed into a
    # pinned batch by the library's host thread pool (PinnedFrameBatch.fill_async -> vn_host_pack_submit), the batch
    # is copied to a device staging slot on a copy stream, the step runs, and its per-frame voxel counts are read
    # back.  The arrays of consecutive steps are distinct objects from a ring larger than the CPU's last-level cache
    # (contents alternate between the two batches like the device-resident leg).  All through the package's public
    # feeding API (dataset.PinnedFrameBatch + pipeline.HostFrameStream).
    from voxelnet_pytorch_b200 import HostFrameStream, PinnedFrameBatch

    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 8)
    n_pack = args.host_threads if args.host_threads else max(2, min(8, cores // max(world, 1) - 1))
    n_pack = _cabi.lib().vn_host_pack_threads(n_pack) or n_pack
    ring_len = 4 if batch <= 16 else 1  # host copies per content: 8 x 31 MB at car16
    frame_sets = (frames, frames_b)
    sets_sorted = sorted(range(2), key=lambda c: -sum(int(f.shape[0]) for f in frame_sets[c]))  # like `batches`
    host_ring = [[[np.array(f, copy=True) for f in frame_sets[c]] for _ in range(ring_len)] for c in sets_sorted]
    max_points = max(sum(int(f.shape[0]) for f in fs) for fs in frame_sets)
    stream = HostFrameStream(fes if P > 1 else fe, batch, max_points, cuda_graphs=not args.no_cuda_graph)
    NP = P + 3
    pinned = [PinnedFrameBatch(batch, max_points) for _ in range(NP)]

    def src(i):
        return host_ring[batch_of(i)][(i // 2) % ring_len]

    # the same ring in page-locked memory: what a loader that reads .bin files into pinned arrays hands over
    # (dataset.load_velodyne_bin(pinned=True)); HostFrameStream.upload_frames copies these straight into the device slot
    from voxelnet_pytorch_b200 import pinned_frame

    def to_pinned(f):
        a = pinned_frame(f.shape[0])
        a[...] = f
        return a

    pinned_ring = [[[to_pinned(f) for f in fs] for fs in ring] for ring in host_ring]

    def src_pinned(i):
        return pinned_ring[batch_of(i)][(i // 2) % ring_len]

    def e2e_run(n, mode, prepacked=None):
        # step i: upload i+P -> forward i -> counts of step i-P.  The host consumes the counts P steps behind (the device
        # queue never drains) and the last P before the clock stops.
        #   mode "pinned_frames": per-frame page-locked arrays, one async copy per frame (no host packing)
        #   mode "pageable":      per-frame pageable arrays, packed into a pinned batch by the host thread pool first
        #   mode "prepacked":     batches that already sit packed in pinned memory (round 1's e2e definition)
        def up(j):
            if mode == "pinned_frames":
                stream.upload_frames(src_pinned(j))
            elif mode == "pageable":
                stream.upload(pinned[j % NP])
            else:
                stream.upload(prepacked[batch_of(j)])

        if mode == "pageable":
            for j in range(min(P + 1, n)):
                pinned[j % NP].fill_async(src(j))
        for j in range(min(P, n)):
            up(j)
        seen = 0
        for i in range(n):
            if mode == "pageable" and i + P + 1 < n:
                pinned[(i + P + 1) % NP].fill_async(src(i + P + 1))
            if i + P < n:
                up(i + P)
            stream.forward()
            if i >= P:
                seen += int(stream.voxel_counts_host(lag=P)[-1])
        for lag in range(min(P, n) - 1, -1, -1):
            seen += int(stream.voxel_counts_host(lag=lag)[-1])
        return seen

    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prepacked = [PinnedFrameBatch.from_frames(frame_sets[c]) for c in sets_sorted]
    ms_by_mode = {}
    for mode in ("pinned_frames", "pageable", "prepacked"):
        e2e_run(4 * P * (P + 1), mode, prepacked)  # (with CUDA graphs: every (front end, staging slot, batch) step is captured on its second visit)
        barrier()
        f0.record()
        e2e_run(args.steps, mode, prepacked)
        f1.record()
        barrier()
        ms_by_mode[mode] = f0.elapsed_time(f1)
    ms_e2e, ms_e2e_pageable, ms_e2e_prepacked = ms_by_mode["pinned_frames"], ms_by_mode["pageable"], ms_by_mode["prepacked"]

    # ---- one step at a time (P = 1) beside the in-flight figure -------------------------------------------
    ms_single = None
    if P > 1:
        for i in range(4):
            (graphs[0][i % 2].replay() if graphs is not None else step(i))
        barrier()
        f0.record()
        for i in range(args.steps):
            (graphs[0][i % 2].replay() if graphs is not None else step(i))
        f1.record()
        barrier()
        ms_single = f0.elapsed_time(f1)

    # ---- the other sparse->dense form beside the timed one (N = 1 only): one step at a time, graph replay ----------
    other_form = None
    if world == 1 and batch <= 16 and graphs is not None:
        fe_o = vn.VoxelNetFrontEnd(fnet, mnet, target, materialise_grid=fe.dense is None)
        g_o = [fe_o.capture_graph(bt.points, bt.offsets, bt.max_n) for bt in batches]
        for i in range(4):
            g_o[i % 2].replay()
        barrier()
        f0.record()
        for i in range(args.steps):
            g_o[i % 2].replay()
        f1.record()
        barrier()
        t_o = StageTimer()
        for i in range(10):
            fe_o.forward_points(batches[i % 2].points, batches[i % 2].offsets, batches[i % 2].max_n, timer=t_o)
        torch.cuda.synchronize()
        other_form = {"form": "resident bf16 grid" if fe_o.dense is not None else "never materialised: row table + occupancy index (vn_conv3d_bn_relu_fwd_rows)",
                      "single_pipeline_ms_per_step": f0.elapsed_time(f1) / args.steps,
                      "stages_ms": {k: v[0] for k, v in t_o.summary().items()},
                      "device_bytes_per_front_end": fe_o.device_bytes(),
                      "note": "bit-identical output (tests/test_gpu_bench_parity.py); compare with single_pipeline.ms_per_step and config.device_bytes_per_front_end"}
        del g_o, fe_o

    # ---- SURVEY 8(f) "next" row: the RPN behind the path (N = 1 only; one step at a time, eager launches) ----------
    # Not part of `value` / `e2e`: the front end hands its map over as bf16 channels-last and the RPN half of the reference's
    # MiddleConvNet (model.py:212-281) runs on the tcgen05 kernel (HybridRPN) - beside the same layers on stock bf16 cuDNN.
    detect = None
    if world == 1 and batch <= 16 and not args.no_rpn:
        try:
            torch.manual_seed(2)
            rpn_ref = vn.RPN(target).to(dev).eval()
            for m in rpn_ref.modules():
                if isinstance(m, torch.nn.BatchNorm2d):
                    m.weight.data = 0.5 + torch.rand(m.weight.shape, generator=g).to(dev)
                    m.bias.data = ((torch.rand(m.bias.shape, generator=g) - 0.5) * 0.6).to(dev)
                    m.running_mean.data = ((torch.rand(m.running_mean.shape, generator=g) - 0.5) * 0.4).to(dev)
                    m.running_var.data = (0.5 + torch.rand(m.running_var.shape, generator=g)).to(dev)
            fe_d = vn.VoxelNetFrontEnd(fnet, mnet, target, output_format="nhwc_bf16")
            n_rep = max(10, min(args.steps, 50))

            def timed(fn):
                for _ in range(3):
                    fn()
                barrier()
                f0.record()
                for i in range(n_rep):
                    fn(i)
                f1.record()
                barrier()
                return f0.elapsed_time(f1) / n_rep

            with torch.no_grad():
                hy = vn.HybridRPN(rpn_ref)
                hy.permute_input_channels(fe_d.reference_channel_of_output())
                x_d = fe_d.forward_points(batches[0].points, batches[0].offsets, batches[0].max_n)
                ms_rpn = timed(lambda i=0: hy(x_d))
                ms_both = timed(lambda i=0: hy(fe_d.forward_points(batches[i % 2].points, batches[i % 2].offsets, batches[i % 2].max_n)))
                st = vn.HybridRPN(rpn_ref)
                st.permute_input_channels(fe_d.reference_channel_of_output())
                st.set_min_tiles(1 << 30)  # every layer on the private bf16 channels-last copies of the stock layers
                ms_stock = timed(lambda i=0: st(x_d))
                prob, reg = hy(x_d)
                p2, r2 = st(x_d)
                torch.cuda.synchronize()
                dev_vs_stock = max(float((prob - p2).abs().max()), float((reg - r2).abs().max() / r2.abs().max().clamp_min(1e-6)))
            detect = {
                "rpn_ms_per_batch": ms_rpn, "rpn_stock_bf16_ms_per_batch": ms_stock,
                "points_to_rpn_maps_ms_per_batch": ms_both, "points_to_rpn_maps_frames_per_s": batch / (ms_both * 1e-3),
                "outputs": [list(prob.shape), list(reg.shape)], "max_dev_vs_stock_bf16": dev_vs_stock, "reps": n_rep,
                "note": "voxelnet_pytorch_b200.HybridRPN(RPN(%r)) behind VoxelNetFrontEnd(output_format='nhwc_bf16'): one step at a time on one "
                        "stream, eager launches, inputs resident; random-init weights.  Outside `value` and `e2e` (SURVEY 8(f) next row); "
                        "parity against the fp32 reference layers: tests/test_gpu_rpn_conv2d.py" % target,
            }
            del hy, st, fe_d, x_d, rpn_ref
        except Exception as e:  # the headline line must still be printed
            detect = {"error": "%s: %s" % (type(e).__name__, e)}

    from voxelnet_pytorch_b200 import sharding

    per_rank_ms = sharding.gather_floats(ms_total / args.steps)
    per_rank_e2e_ms = sharding.gather_floats(ms_e2e / args.steps)
    ms_total, ms_e2e, ms_e2e_prepacked, ms_e2e_pageable = sharding.reduce_max([ms_total, ms_e2e, ms_e2e_prepacked, ms_e2e_pageable])  # device time, max over ranks
    if ms_single is not None:
        ms_single = sharding.reduce_max([ms_single])[0]
    frames_total = sum(sharding.gather_ints(batch))

    stages = {k: v[0] for k, v in timer.summary().items()}

    if rank == 0:
        peaks = load_peaks()
        c = fe.cfg
        flops = conv_flops_per_frame(c.depth, c.height, c.width)
        ms_step = ms_total / args.steps
        value = frames_total / (ms_step * 1e-3)
        K_total = sum(cc[-1] for cc in counts) / len(counts)
        N_total = sum(int(bt.offs_np[-1]) for bt in batches) / 2
        T = c.points_per_voxel
        # SURVEY 8(d) algorithmic bytes (the materialised data flow: [K,T,7] written then read) ...
        bytes_vox = 16 * N_total + 28 * T * K_total + 32 * K_total + 8 * K_total
        bytes_svfe = 28 * T * K_total + 512 * K_total
        # ... and what the fused kernels really move: the feature buffer never exists
        moved_vox = 16 * N_total + 5 * 4 * N_total                      # points in, per-point key/row/slot/bucket words
        moved_svfe = 16 * N_total / 2 + 4 * N_total / 2 + 512 * K_total + 40 * K_total
        bytes_scatter = 512 * K_total + 256 * K_total + 32 * K_total
        stage_roofline = {
            "voxelize": {"bound": "hbm", "achieved": bytes_vox / (stages["voxelize"] * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "moved_gbs": moved_vox / (stages["voxelize"] * 1e-3) / 1e9},
            "svfe": {"bound": "hbm", "achieved": bytes_svfe / (stages["svfe"] * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "moved_gbs": moved_svfe / (stages["svfe"] * 1e-3) / 1e9,
                     "fp32_tflops": (N_total / 2 + K_total) * (2 * 7 * 16 + 2 * 32 * 64) / (stages["svfe"] * 1e-3) / 1e12},
            "scatter": {"bound": "hbm", "achieved": bytes_scatter / (stages["scatter"] * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s"},
        }
        # Conv layers.  The kernels skip input boxes no voxel can reach and store only output blocks that
        # are fed now or were fed in the previous step, so two figures are given per layer:
        #   executed   FLOPs of the tcgen05.mma instructions actually issued (from the device-built plan)
        #   dense_eq   SURVEY 8(d)'s dense-equivalent FLOPs of the layer / time (can exceed the peak)
        # and the HBM side: unique live input blocks + stored output blocks + the weights once.
        out_bytes = [128 * 64 * 2, 128 * 64 * 2, 128 * 64 * (4 if args.output == "nchw_f32" else 2)]
        wbytes = [27 * 128 * 64 * 2, 27 * 64 * 64 * 2, 27 * 64 * 64 * 2]
        conv_roof = []
        for l, name in enumerate(("conv1", "conv2", "conv3")):
            t = stages[name] * 1e-3
            st = plan[l]
            hbm = st["unique_input_bytes"] + (st["touched_blocks"] + st["restored_blocks"]) * out_bytes[l] + wbytes[l]
            ent = {
                "tensor": {"achieved": st["executed_flops"] / t / 1e12, "peak": peaks["tflops"], "unit": "TFLOP/s"},
                "hbm": {"achieved": hbm / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s", "bytes_per_launch": hbm},
                "dense_equivalent_tflops": flops[l] * batch / t / 1e12,
                "live_box_fraction": st["live_boxes"] / st["boxes_total"],
                "stored_block_fraction": (st["touched_blocks"] + st["restored_blocks"]) / st["blocks_total"],
                "restored_block_fraction": st["restored_blocks"] / st["blocks_total"],
                "groups": st["groups"],
            }
            for k in ("tensor", "hbm"):
                ent[k]["frac"] = ent[k]["achieved"] / ent[k]["peak"]
            ent["bound"] = "tensor" if ent["tensor"]["frac"] >= ent["hbm"]["frac"] else "hbm"
            stage_roofline[name] = ent
            conv_roof.append(ent)
        for k in ("voxelize", "svfe", "scatter"):
            stage_roofline[k]["frac"] = stage_roofline[k]["achieved"] / stage_roofline[k]["peak"]
        traffic = None
        try:  # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture
            with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
                traffic_tab = json.load(f)
        except Exception:
            traffic_tab = {}
        dom = max(range(3), key=lambda l: stages["conv%d" % (l + 1)])
        if world == 1 and args.config in traffic_tab:
            traffic = traffic_tab[args.config].get("conv%d" % (dom + 1))
        dk = conv_roof[dom]
        kind = dk["bound"]
        line = {
            "metric": metric_name(args.config), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if args.config == "stress64" else "weak", "vs_baseline": None,
            "dtype": "bf16", "data": DATA,
            "config": {"workload": desc, "name": args.config, "batch_per_gpu": batch, "points_per_batch": N_total,
                       "voxels_per_batch": K_total,
                       "inputs": "two different batches alternate step by step (resident outputs are restored where the previous batch reached)",
                       "l2_policy": "working set (%.2f GB of activations, outputs and row tables per front end) exceeds L2" % (fe.device_bytes()["total"] / 1e9),
                       "device_bytes_per_front_end": fe.device_bytes(),
                       "grid": "never materialised: (K,128) bf16 rows + the voxelizer's occupancy index feed conv-1 (vn_conv3d_bn_relu_fwd_rows)" if fe.dense is None
                               else "resident bf16 grid (default policy: it fits the 16 GiB budget); --grid rows times the never-materialised form",
                       "vfe": "fp32", "conv": "bf16 operands, fp32 accumulate (tcgen05)",
                       "output": "float32 (B,128,H,W), the reference's tensor" if args.output == "nchw_f32" else "bf16 channels-last (B,128,H,W) view, channel d*64+c (option for a bf16 RPN; NOT the drop-in tensor)",
                       "launch": launch_mode,
                       "in_flight": "%d step(s) in flight per GPU (%s); ms_per_step = timed region / steps" % (
                           P, "one front end per pipeline stream, own grid/workspaces/outputs" if P > 1 else "one stream"),
                       "stage_times": "stages_ms / stage_roofline come from a separate eager pass of ONE front end with CUDA events between the stages; with several steps in flight their sum exceeds ms_per_step (the steps overlap)"},
            "ms_per_frame": ms_step / batch,
            "stages_ms": stages,
            "clocks": clocks,
            "e2e": {"value": frames_total / (ms_e2e / args.steps * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(sum(bt.host_points.numel() * 4 + bt.host_offsets.numel() * 4 for bt in batches) / 2),
                    "d2h_bytes_per_step": int(host_counts.numel() * 4),
                    "from": "a list of per-frame (N_i,4) float32 numpy arrays per step (what the reference's loader yields, dataset.py:51) held in "
                            "page-locked memory (dataset.load_velodyne_bin(pinned=True) / pinned_frame): HostFrameStream.upload_frames issues one "
                            "asynchronous copy per frame into the device CSR slot, no host-side packing; fresh array objects every step from a "
                            "ring of %d x %.0f MB" % (2 * ring_len, max_points * 16 / 1e6),
                    "value_from_pageable_arrays": frames_total / (ms_e2e_pageable / args.steps * 1e-3),
                    "pageable_note": "the same list in ordinary pageable memory: packed into a pinned batch by the library's host thread pool "
                                     "(%d threads per rank) before the copy - host-memory-bound when 8 ranks share one host" % int(n_pack),
                    "host_pack_threads": int(n_pack),
                    "value_from_prepacked_pinned_batches": frames_total / (ms_e2e_prepacked / args.steps * 1e-3),
                    "ms_per_step_by_rank": per_rank_e2e_ms,
                    "note": "every step inside the timed region: %d per-frame H2D copies on a copy stream, the step "
                            "(voxelnet_pytorch_b200.HostFrameStream, %d front end(s) in flight, %d staging buffers), D2H of its per-frame voxel counts "
                            "(consumed %d step(s) later, the last ones before the clock stops).  The %.2f GB (B,128,H,W) result stays on the device: its consumer, "
                            "the RPN, runs on the same GPU" % (batch, P, P + 1, P, fe.out.numel() * 4 / 1e9)},
            "ms_per_step_by_rank": per_rank_ms,
            "single_pipeline": None if ms_single is None else {
                "value": frames_total / (ms_single / args.steps * 1e-3), "ms_per_step": ms_single / args.steps,
                "note": "ONE step at a time (graph replays back to back on one stream): the latency of a step; `value` above is throughput with %d steps in flight" % P},
            "other_sparse_to_dense_form": other_form,
            "detection_pipeline": detect,
            "gpu_launches": int(launches),
            "roofline": {"kernel": "k_conv3d_bn_relu (layer %d)" % (dom + 1), "bound": kind, "achieved": dk[kind]["achieved"],
                         "peak": dk[kind]["peak"], "unit": dk[kind]["unit"], "frac": dk[kind]["frac"], "traffic": traffic,
                         "traffic_source": traffic_tab.get("source") if traffic is not None else None,
                         "peak_source": peaks["source"] + (" (bf16_tflops_sustained: cuBLAS back to back for seconds, power-limited; a kernel timed inside this step can exceed it - "
                                                           "the burst figure is %.0f TFLOP/s, frac_of_burst below)" % peaks["tflops_burst"] if kind == "tensor" else " (hbm_gbs)"),
                         "frac_of_burst": dk[kind]["achieved"] / peaks["tflops_burst"] if kind == "tensor" else None,
                         "counts": "executed work: MMAs issued / blocks moved, decoded from the device-built plan; "
                                   "dense-equivalent %.0f TFLOP/s" % dk["dense_equivalent_tflops"],
                         "other_bound": {k: dk[k] for k in ("tensor", "hbm") if k != kind}},
            "stage_roofline": stage_roofline,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import model_torch as OM

            cpu = cpu_reference_sample(target, frames[:4], OM.random_state_dict(0), args.cpu_seconds)
            line["cpu_baseline"] = {
                "value": 1.0 / cpu["best_s"], "unit": UNIT, "cores": cpu["cores"], "kind": cpu["kind"],
                "sample": "%d single frames of the same workload (best of): %s" % (cpu["frames"], cpu["note"]),
                "mean_s_per_frame": cpu["mean_s"], "stages_mean_s": cpu["stages_mean_s"], "host_cpus": cpu["host_cpus"],
            }
        emit(line)

    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
