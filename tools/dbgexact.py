import ctypes, torch, sys, time
sys.path.insert(0,'.')
from data_compression_b200 import api
from data_compression_b200.fields import make_field_torch
dev=torch.device('cuda:0'); n=1<<28
x=make_field_torch('smooth', n, dev)
ws=api.Workspace(n,4096,dev)
for i in range(3):
    comp=api.compress(x, ct=7, abs_error_bound=0.01, chunk_elems=4096, workspace=ws)
torch.cuda.synchronize()
L=api.lib(); buf=(ctypes.c_ulonglong*4)()
L.dcb_debug_exact(buf)
v=list(buf); print('exact_medium: %.0f cycles per call (%.2f us at 1.92 GHz), calls %d (per launch %d)'%(v[0]/max(v[1],1), v[0]/max(v[1],1)/1920, v[1], v[1]//3))
