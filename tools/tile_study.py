"""Executed vs useful FLOPs of the three middle layers under different output-tile shapes (CPU, numpy/torch).

Model = what k_conv3d_bn_relu executes: a box (tile, input plane p, kw) is live when any cell of the layer's input that can
deviate from its background lies in rows [TH*th-1, TH*th+TH] x cols [TW*tw+kw-1, TW*tw+kw+TW-2] of plane p; a live box costs
nkd(p) * 3 (kh) * TH*TW * 64 (k) * 64 (n) * 2 FLOPs per 64-channel chunk.  Useful = taps that really hit a deviating cell.
Usage: python tools/tile_study.py [kitti|synth]
"""
import sys

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, '/root/repo')
from oracle import voxelize_np as OV

LAYERS = [(128, 2, 1), (64, 1, 0), (64, 2, 1)]  # cin, stride_d, pad_d


def planes(din, stride, pad):
    dout = (din + 2 * pad - 3) // stride + 1
    nkd = []
    for p in range(din):
        n = 0
        for kd in range(3):
            num = p + pad - kd
            if num >= 0 and num % stride == 0 and num // stride < dout:
                n += 1
        nkd.append(n)
    return dout, nkd


def reach(occ, stride, pad, dout):
    """occ [D,H,W] bool -> outputs [Dout,H,W] that see at least one active input."""
    D, H, W = occ.shape
    x = torch.from_numpy(occ.astype(np.float32))[None, None]
    y = F.max_pool3d(F.pad(x, (1, 1, 1, 1, pad, pad)), kernel_size=3, stride=(stride, 1, 1))
    return y[0, 0].numpy() > 0


def study(frames, shapes):
    tot = {s: [0.0, 0.0, 0.0] for s in shapes}
    useful = [0.0, 0.0, 0.0]
    dense = [0.0, 0.0, 0.0]
    for f in frames:
        v = OV.pcl_to_voxels_fast(f, 'Car')
        occ = np.zeros((10, 400, 352), bool)
        c = v['coordinate_buffer']
        occ[c[:, 0], c[:, 1], c[:, 2]] = True
        din = 10
        for l, (cin, stride, pad) in enumerate(LAYERS):
            dout, nkd = planes(din, stride, pad)
            H, W = occ.shape[1:]
            # useful: (active input cell, tap) pairs that land on an existing output
            per_plane = occ.reshape(din, -1).sum(1)
            useful[l] += sum(per_plane[p] * nkd[p] for p in range(din)) * 9 * cin * 64 * 2
            dense[l] += dout * H * W * 27 * cin * 64 * 2
            for (TH, TW) in shapes:
                th_n, tw_n = -(-H // TH), -(-W // TW)
                ex = 0.0
                o = torch.from_numpy(occ.astype(np.float32))  # [D,H,W]
                # rows: any active in [TH*th-1, TH*th+TH]
                op = F.pad(o, (0, 0, 1, TH * th_n - H + 1))  # pad H by 1 top, fill bottom
                rows = F.max_pool2d(op[None], kernel_size=(TH + 2, 1), stride=(TH, 1))[0]  # [D, th_n, W]
                for kw in range(3):
                    # cols [TW*tw+kw-1, TW*tw+kw+TW-2]
                    rp = F.pad(rows, (1, TW * tw_n - W + 1))
                    cols = F.max_pool1d(rp[:, :, kw:kw + TW * tw_n].reshape(din * th_n, 1, -1)[:, 0][None], kernel_size=TW, stride=TW)[0]
                    live = cols.reshape(din, th_n, tw_n) > 0
                    for p in range(din):
                        ex += float(live[p].sum()) * nkd[p] * 3 * TH * TW * 64 * 64 * 2 * (cin // 64)
                tot[(TH, TW)][l] += ex
            occ = reach(occ, stride, pad, dout)
            din = dout
    return tot, useful, dense


if __name__ == '__main__':
    kind = sys.argv[1] if len(sys.argv) > 1 else 'synth'
    if kind == 'kitti':
        from oracle import ref_import as R
        frames = [np.fromfile(R.kitti_frame_path(i), dtype=np.float32).reshape(-1, 4) for i in (0, 10, 20, 31)]
    else:
        from voxelnet_pytorch_b200 import synth
        frames = synth.lidar_batch(4, seed0=0)
    shapes = [(16, 8), (8, 16), (8, 8), (4, 16), (16, 4), (4, 8), (32, 4), (4, 4), (2, 8), (1, 8)]
    tot, useful, dense = study(frames, shapes)
    n = len(frames)
    print(kind, 'GFLOP per frame: dense', [round(d / n / 1e9, 1) for d in dense], 'useful', [round(u / n / 1e9, 1) for u in useful])
    for s in shapes:
        print('tile %2dx%-2d executed GFLOP/frame' % s, [round(t / n / 1e9, 1) for t in tot[s]], 'sum %.1f' % (sum(tot[s]) / n / 1e9),
              ' x useful', [round(t / u, 2) for t, u in zip(tot[s], useful)])
