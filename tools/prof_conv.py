import sys, ctypes as C, json
sys.path.insert(0,'/root/repo')
import torch, numpy as np
import voxelnet_pytorch_b200 as vn
from voxelnet_pytorch_b200 import _cabi, synth
dev=torch.device('cuda',0)
torch.manual_seed(0)
fnet=vn.FeatureLearningNet('Car').eval(); mnet=vn.MiddleConvNet().eval()
fe=vn.VoxelNetFrontEnd(fnet.to(dev),mnet.to(dev),'Car')
A=synth.lidar_batch(16,seed0=0); B=synth.lidar_batch(16,seed0=500)
pa=vn.pack_frames(A); pb=vn.pack_frames(B)
for i in range(4):
    p=pa if i%2==0 else pb
    fe.forward_points(*p)
torch.cuda.synchronize()
lib=_cabi.lib()
lib.vn_conv3d_debug_profile.argtypes=[C.c_void_p]; lib.vn_conv3d_debug_profile.restype=None
# profile each conv separately: the buffer is overwritten by each conv launch, so run one step per layer with hooks
from voxelnet_pytorch_b200 import ops
orig=ops.conv3d_bn_relu_fwd
bufs=[]
def wrapped(*a,**k):
    buf=torch.zeros(148*16,dtype=torch.int64,device=dev)
    lib.vn_conv3d_debug_profile(C.c_void_p(buf.data_ptr()))
    r=orig(*a,**k)
    lib.vn_conv3d_debug_profile(None)
    bufs.append(buf)
    return r
ops.conv3d_bn_relu_fwd=wrapped
import voxelnet_pytorch_b200.pipeline as pl
fe.forward_points(*pa)
torch.cuda.synchronize()
names=['p_total','p_wait_empty_a','p_wait_empty_b','p_boxes','m_total','m_wait_full_a','m_wait_full_b','m_wait_acc_empty','e_total','e_wait_sched','e_wait_acc_full']
for l,b in enumerate(bufs):
    x=b.cpu().numpy().reshape(148,16)
    print('conv%d'%(l+1), {n:int(x[:,i].mean()) for i,n in enumerate(names)}, 'max total', int(x[:,4].max()), 'min', int(x[:,4].min()))
