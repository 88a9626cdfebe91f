"""Execution context: the stand-in for `SparkSession.builder().master("local[*]").getOrCreate()`
(DataQuality4MachineLearningApp.java:38-41) -- one object per process that owns the GPU resources of the path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi


class Context:
    """One CUDA device + stream + scratch (dq4ml_ctx).  Not thread-safe: one per thread, like a Spark task."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        self._cb = None
        capi.check(capi.lib().dq4ml_ctx_create(int(device), C.byref(self._h)))
        self.device = int(device)
        self.rank, self.world_size = 0, 1

    def close(self):
        if self._h:
            capi.lib().dq4ml_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def synchronize(self):
        capi.check(capi.lib().dq4ml_ctx_synchronize(self._h))

    @property
    def stream(self) -> int:
        return int(capi.lib().dq4ml_ctx_stream(self._h) or 0)

    @property
    def launch_count(self) -> int:
        return int(capi.lib().dq4ml_ctx_launch_count(self._h))

    def set_gram_tier(self, tier="fp64"):
        """"fp64" (default, 1e-9 parity) or "tf32x3" (tcgen05 tensor cores, TMEM accumulators, 1e-5 tolerance, d <= 256)."""
        tiers = {"fp64": 0, "tf32x3": 1}
        if tier not in tiers:
            raise ValueError(f"tier must be one of {sorted(tiers)}")
        capi.check(capi.lib().dq4ml_ctx_set_gram_tier(self._h, tiers[tier]))

    def enable_pass_timing(self, on=True):
        capi.check(capi.lib().dq4ml_ctx_enable_pass_timing(self._h, int(bool(on))))

    def pass_timing(self):
        """(summed kernel ms, number of aggregation passes) since enable_pass_timing."""
        ms, cnt = C.c_double(0.0), C.c_int64(0)
        capi.check(capi.lib().dq4ml_ctx_pass_timing(self._h, C.byref(ms), C.byref(cnt)))
        return float(ms.value), int(cnt.value)

    def pass_times(self):
        """Kernel time (ms) of every aggregation pass since enable_pass_timing, in launch order."""
        buf = np.zeros(4096, dtype=np.float64)
        n = capi.check(capi.lib().dq4ml_ctx_pass_times(self._h, buf.ctypes.data, buf.size))
        return buf[:min(n, buf.size)].tolist()

    # ---- ranks -------------------------------------------------------------------------------------------------
    def bind_torch_distributed(self, use_nccl=None):
        """Row-sharded mode: binds this ctx to the torch.distributed world (one process per GPU).

        With the nccl backend the library opens its own NCCL communicator (unique id broadcast through
        torch.distributed) and sums result vectors on-device; with gloo (CPU tests) a host callback is used.
        """
        import torch
        import torch.distributed as dist

        if not dist.is_initialized():
            return self
        self.rank, self.world_size = dist.get_rank(), dist.get_world_size()
        if self.world_size == 1:
            return self
        if use_nccl is None:
            use_nccl = dist.get_backend() == "nccl"
        if use_nccl:
            ident = np.zeros(128, dtype=np.uint8)
            if self.rank == 0:
                capi.check(capi.lib().dq4ml_nccl_unique_id(ident.ctypes.data))
            t = torch.from_numpy(ident).to(f"cuda:{self.device}")
            dist.broadcast(t, src=0)
            ident = t.cpu().numpy()
            capi.check(capi.lib().dq4ml_ctx_init_nccl(self._h, ident.ctypes.data, self.rank, self.world_size))
        else:
            self.set_allreduce(torch_allreduce_callback())
        return self

    def set_allreduce(self, fn, rank=None, world_size=None):
        """fn(numpy float64 array) sums it in place over ranks; rank / world_size (default: this Context's, as set by
        bind_torch_distributed) tell the library which per-rank slot of the min / max and centre vectors is its own."""
        if rank is not None:
            self.rank = int(rank)
        if world_size is not None:
            self.world_size = int(world_size)

        def _cb(_user, ptr, count):
            try:
                a = np.ctypeslib.as_array(ptr, shape=(count,))
                fn(a)
                return 0
            except Exception:  # never let an exception cross the C boundary
                return 1

        self._cb = capi.ALLREDUCE_FN(_cb)
        capi.check(capi.lib().dq4ml_ctx_set_allreduce(self._h, self._cb, None, self.rank, self.world_size))

    def allreduce(self, a: np.ndarray):
        a = np.ascontiguousarray(a, dtype=np.float64)
        capi.check(capi.lib().dq4ml_ctx_allreduce(self._h, a.ctypes.data, a.size))
        return a

    # ---- synthetic data --------------------------------------------------------------------------------------
    def synth_generate(self, n, d=1, row0=0, seed_data=20200601, seed_anom=7, p_low=0.01, p_high=0.01):
        """Device-resident benchmark columns: guest:int32[n], price:f64[n], extra:f64[d-1, n] (torch cuda tensors)."""
        import torch

        dev = f"cuda:{self.device}"
        guest = torch.empty(n, dtype=torch.int32, device=dev)
        price = torch.empty(n, dtype=torch.float64, device=dev)
        stride = (n + 15) // 16 * 16  # keeps every column 16-byte aligned whatever n is
        store = torch.empty((max(d - 1, 0), stride), dtype=torch.float64, device=dev)
        sp = capi.SynthParams(seed_data, seed_anom, p_low, p_high, d)
        capi.check(capi.lib().dq4ml_synth_generate(self._h, C.byref(sp), row0, n, guest.data_ptr(), price.data_ptr(),
                                                    store.data_ptr() if d > 1 else None, stride))
        extra = store[:, :n]  # extra[k] is a contiguous 1-D column
        return guest, price, extra

    def read_csv(self, data: bytes, to_device=True):
        """spark.read().format("csv").option("inferSchema","true").option("header","false").load(...) for the lab's
        two-column dialect (DataQuality4MachineLearningApp.java:52-59): returns (guest:int32, price:f64), parsed on the
        GPU, as torch cuda tensors (or numpy arrays when to_device is False)."""
        import torch

        n_bytes = len(data)
        cap = n_bytes // 4 + 1  # the shortest record, "1,2" + separator, is 4 bytes
        if to_device:
            dev = f"cuda:{self.device}"
            guest = torch.empty(cap, dtype=torch.int32, device=dev)
            price = torch.empty(cap, dtype=torch.float64, device=dev)
        else:
            guest = np.empty(cap, dtype=np.int32)
            price = np.empty(cap, dtype=np.float64)
        n = C.c_int64(0)
        buf = C.create_string_buffer(data, n_bytes) if n_bytes else None
        capi.check(capi.lib().dq4ml_csv_parse(self._h, C.cast(buf, C.c_void_p) if buf is not None else None, n_bytes,
                                               capi.buf_ptr(guest), capi.buf_ptr(price), cap, C.byref(n)))
        m = int(n.value)
        if to_device:
            return guest[:m].clone(), price[:m].clone()  # own, aligned storage of the right length
        return guest[:m].copy(), price[:m].copy()

    def rows_digest(self, guest, price, seed=0):
        out = np.zeros(2, dtype=np.uint64)
        capi.check(capi.lib().dq4ml_rows_digest(self._h, capi.buf_ptr(guest), capi.buf_ptr(price), capi.buf_len(guest),
                                                 seed, out.ctypes.data))
        return int(out[0]), int(out[1])


def torch_allreduce_callback():
    import torch
    import torch.distributed as dist

    def fn(a: np.ndarray):
        t = torch.from_numpy(a)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)

    return fn


def shard_bounds(n: int, rank: int, world: int, align: int = 1024):
    """Contiguous row range [lo, hi) of `rank`; boundaries are multiples of `align` rows so that every shard of a
    16-byte aligned column is itself 16-byte aligned (partition r gets rows [r n/G, (r+1) n/G))."""
    def cut(r):
        if r >= world:
            return n
        return min(n, (n * r // world) // align * align)
    return cut(rank), cut(rank + 1)
