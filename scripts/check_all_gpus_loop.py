import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sctl_b200 import KERNELS, capi, synth
d = capi.kernel_dims(3)
Xt, Xs, Xn, F = synth.make_problem(d, 20000, 9000, np.float64)
ker = KERNELS[3]()
U1 = ker.Eval(None, Xt, Xs, Xn, F)
import torch
free0 = [torch.cuda.mem_get_info(i)[0] for i in range(torch.cuda.device_count())]
for i in range(30):
    U = ker.Eval(None, Xt, Xs, Xn, F, all_gpus=True)
    assert np.allclose(U, U1, rtol=1e-12, atol=1e-14), i
free1 = [torch.cuda.mem_get_info(i)[0] for i in range(torch.cuda.device_count())]
print("ALL_GPUS x30 OK on", capi.device_count(), "devices; free memory delta (MB):", [(a - b) / 1e6 for a, b in zip(free0, free1)])
