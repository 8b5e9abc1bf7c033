"""Per-stage timeline of the chain kernel (diagnostic)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from tensorflow_train_distributed_b200.config import TrainConfig
from tensorflow_train_distributed_b200.ops import _native as N
from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer

cfg = TrainConfig(worker_hosts="127.0.0.1:2222", save_checkpoint_secs=None, cuda_graph=False, step_impl="layers")
tr = GpuTrainer(cfg)
tr.initialize()
lib = N.load()
n = tr.model._chain.n
buf = torch.zeros(4 * n, dtype=torch.int64, device="cuda")
lib.ttd_gemm_chain_set_trace.argtypes = [C.c_void_p]
lib.ttd_gemm_chain_set_trace(buf.data_ptr())
x = torch.rand(128, 784); y = torch.randint(0, 10, (128,))
for _ in range(3):
    tr.train_step(x, y)
t = buf.cpu().view(n, 4)
t0 = int(t[0, 0])
for s in range(n):
    a, b, c, d = (int(v) - t0 for v in t[s])
    print(f"stage {s:2d}: start {a/1e3:8.2f} us | operands +{(b-a)/1e3:6.2f} | mma +{(c-b)/1e3:6.2f} | epilogue +{(d-c)/1e3:6.2f}")
