import os, sys, time, faulthandler
faulthandler.dump_traceback_later(40, exit=True)
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import sparkdq4ml_b200 as pkg
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
def log(*a): print(f"[r{rank} {time.time()%1000:.2f}]", *a, flush=True)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
log("pg up")
t = torch.ones(4, device="cuda"); dist.all_reduce(t); torch.cuda.synchronize(); log("torch allreduce ok", t[0].item())
ctx = pkg.Context(local); log("ctx")
ident = np.zeros(128, dtype=np.uint8)
if rank == 0:
    pkg.capi.check(pkg.capi.lib().dq4ml_nccl_unique_id(ident.ctypes.data)); log("uid", ident[:8])
tt = torch.from_numpy(ident).cuda(); dist.broadcast(tt, src=0); ident = tt.cpu().numpy(); log("uid bcast", ident[:8])
rc = pkg.capi.lib().dq4ml_ctx_init_nccl(ctx.handle, ident.ctypes.data, rank, world); log("init rc", rc, pkg.capi.lib().dq4ml_last_error())
a = np.arange(6, dtype=np.float64) + rank
ctx.rank, ctx.world_size = rank, world
out = ctx.allreduce(a); log("allreduce", out)
g, p, _ = ctx.synth_generate(1_000_000, d=1, row0=rank * 1_000_000)
from sparkdq4ml_b200 import ml
m = ml.wls_moments(ctx, [g], p, dq=(g, p)); log("moments", m[:4])
dist.barrier(); log("barrier")
dist.destroy_process_group(); log("done")
