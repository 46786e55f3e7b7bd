import ctypes, torch, sys
sys.path.insert(0,'/root/repo')
from data_compression_b200 import api
from data_compression_b200.fields import make_field_torch
dev=torch.device('cuda:0'); n=1<<28
x=make_field_torch('smooth', n, dev)
comp=api.compress(x, ct=7, abs_error_bound=0.01, chunk_elems=4096)
torch.cuda.synchronize()
L=api.lib(); buf=(ctypes.c_ulonglong*8)()
L.dcb_debug_walk(buf)
v=list(buf); print('level2 cycles/tile %.0f  walk loop cycles/tile %.0f  tiles %d  stops/tile %.1f'%(v[0]/v[2], v[1]/v[2], v[2], v[3]/v[2]))
