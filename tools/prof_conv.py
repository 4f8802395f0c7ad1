"""Where does the conv kernel's time go?  (GPU box; debug instantiation of k_conv3d_bn_relu.)

For every experiment switch of `vn_conv3d_debug_flags` (no input-box loads / no weight loads / no MMAs / no stores, and
combinations) the three conv layers of a car16 step are timed with CUDA events; the first row (flags 0, instrumented
kernel) also prints the in-kernel cycle counters.  Results of the switched runs are wrong by construction - only the
times matter.  Usage: python tools/prof_conv.py [out.json]
"""
import ctypes as C
import json
import sys

sys.path.insert(0, '/root/repo')
import torch

import voxelnet_pytorch_b200 as vn
from voxelnet_pytorch_b200 import _cabi, ops, synth
from voxelnet_pytorch_b200.pipeline import StageTimer

dev = torch.device('cuda', 0)
torch.manual_seed(0)
fnet = vn.FeatureLearningNet('Car').eval()
mnet = vn.MiddleConvNet().eval()
fe = vn.VoxelNetFrontEnd(fnet.to(dev), mnet.to(dev), 'Car')
A = synth.lidar_batch(16, seed0=0)
B = synth.lidar_batch(16, seed0=500)
pa, pb = vn.pack_frames(A), vn.pack_frames(B)
if pa[0].shape[0] < pb[0].shape[0]:
    pa, pb = pb, pa
for i in range(4):
    fe.forward_points(*(pa if i % 2 == 0 else pb))
torch.cuda.synchronize()
lib = _cabi.lib()
lib.vn_conv3d_debug_profile.argtypes = [C.c_void_p]
lib.vn_conv3d_debug_profile.restype = None
lib.vn_conv3d_debug_flags.argtypes = [C.c_int32]
lib.vn_conv3d_debug_flags.restype = None

names = ['p_total', 'p_wait_empty_a', 'p_wait_empty_b', 'p_boxes', 'm_total', 'm_wait_full_a', 'm_wait_full_b',
         'm_wait_acc_empty', 'e_total', 'e_wait_sched', 'e_wait_acc_full', 'f0_wait_request', 'f0_lookup', 'f0_wait_stage',
         'f0_clean_store', 'f0_fence']
out = {}


def timed(flags, reps=6):
    lib.vn_conv3d_debug_flags(flags)
    t = StageTimer()
    for i in range(reps):
        fe.forward_points(*(pa if i % 2 == 0 else pb), timer=t)
    torch.cuda.synchronize()
    lib.vn_conv3d_debug_flags(0)
    return {k: round(v[0], 4) for k, v in t.summary().items()}


base = timed(0)
print('production kernel      ', {k: base[k] for k in ('conv1', 'conv2', 'conv3')})
out['production'] = base
LABEL = {16: 'no fill fence', 32: 'no lookups', 48: 'no lookups, no fence', 0: 'instrumented', 1: 'no A loads', 2: 'no B loads', 3: 'no A+B loads', 4: 'no MMA', 8: 'no stores', 11: 'no A+B loads, no stores',
         12: 'no MMA, no stores', 7: 'no loads, no MMA', 15: 'nothing but control'}
# flags 0 with the profile buffer = instrumented kernel
orig = ops.conv3d_bn_relu_fwd
orig_rows = ops.conv3d_bn_relu_fwd_rows
bufs = []


def wrap(fn):
    def wrapped(*a, **k):
        buf = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
        lib.vn_conv3d_debug_profile(C.c_void_p(buf.data_ptr()))
        r = fn(*a, **k)
        lib.vn_conv3d_debug_profile(None)
        bufs.append(buf)
        return r
    return wrapped


ops.conv3d_bn_relu_fwd, ops.conv3d_bn_relu_fwd_rows = wrap(orig), wrap(orig_rows)
t = StageTimer()
fe.forward_points(*pa, timer=t)
torch.cuda.synchronize()
ops.conv3d_bn_relu_fwd, ops.conv3d_bn_relu_fwd_rows = orig, orig_rows
out['counters'] = {}
for l, b in enumerate(bufs):
    x = b.cpu().numpy().reshape(148, 16)
    row = {n: int(x[:, i].mean()) for i, n in enumerate(names)}
    row['m_total_max'], row['m_total_min'] = int(x[:, 4].max()), int(x[:, 4].min())
    out['counters']['conv%d' % (l + 1)] = row
    print('conv%d counters' % (l + 1), row)
for flags in (16, 32, 48, 1, 2, 3, 8, 11, 4, 12, 7, 15):
    r = timed(flags)
    out[LABEL[flags]] = r
    print('%-24s' % LABEL[flags], {k: r[k] for k in ('conv1', 'conv2', 'conv3')})
# ---- per-message trace of CTA 0 (conv-1 .. conv-3 of one step) ------------------------------------------
lib.vn_conv3d_debug_trace.argtypes = [C.c_void_p, C.c_int32]
lib.vn_conv3d_debug_trace.restype = None
import numpy as np

NTR = 1024
traces = []


def trace(fn):
    def traced(*a, **k):
        tb = torch.zeros(NTR * 4, dtype=torch.int64, device=dev)
        pbuf = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
        lib.vn_conv3d_debug_profile(C.c_void_p(pbuf.data_ptr()))
        lib.vn_conv3d_debug_trace(C.c_void_p(tb.data_ptr()), NTR)
        r = fn(*a, **k)
        lib.vn_conv3d_debug_trace(None, 0)
        lib.vn_conv3d_debug_profile(None)
        traces.append(tb)
        return r
    return traced


for flags in (0, 15):
    traces.clear()
    ops.conv3d_bn_relu_fwd, ops.conv3d_bn_relu_fwd_rows = trace(orig), trace(orig_rows)
    lib.vn_conv3d_debug_flags(flags)
    fe.forward_points(*pa, timer=StageTimer())
    torch.cuda.synchronize()
    lib.vn_conv3d_debug_flags(0)
    ops.conv3d_bn_relu_fwd, ops.conv3d_bn_relu_fwd_rows = orig, orig_rows
    for l, tb in enumerate(traces):
        x = tb.cpu().numpy().reshape(NTR, 4)
        box = (x[:, 0] != 0) & (x[:, 1] != 0)
        idx = np.nonzero(box)[0]
        if len(idx) < 16:
            continue
        t_p, t_f, t_i = x[idx, 0], x[idx, 1], x[idx, 2]
        # stage reuse: message j reuses the stage of message j-4 (control messages share the ring)
        reuse = [(x[j, 0] - x[j - 4, 2]) for j in idx if j >= 4 and x[j - 4, 2] != 0]
        row = {
            'messages': int(len(idx)),
            'producer_to_issuer_clk_median': float(np.median(t_f - t_p)),
            'issuer_busy_per_msg_clk_median': float(np.median(t_i - t_f)),
            'issue_done_to_stage_reacquired_clk_median': float(np.median(reuse)) if reuse else None,
            'producer_acquire_interval_clk_median': float(np.median(np.diff(t_p))),
            'producer_acquire_interval_clk_mean': float(np.mean(np.diff(t_p))),
            'issuer_seen_interval_clk_mean': float(np.mean(np.diff(t_f))),
        }
        out['trace_flags%d_conv%d' % (flags, l + 1)] = row
        print('trace flags=%d conv%d' % (flags, l + 1), row)
        if flags == 0 and l == 0:
            out['trace_raw_conv1'] = x[:400].tolist()
stats = fe.plan_stats()
out['plan'] = stats
print(stats)
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], 'w'), indent=1)
