"""b200pt_film_reduce on N GPUs: every rank renders its stripe of sample indices of a small frame, the films are summed
onto rank 0 by the library's own ncclReduce (a communicator made here through ctypes, as a C++ host would with
ncclCommInitRank - torch.distributed only carries the unique id), and rank 0 compares with the whole frame rendered
on its own GPU.   torchrun --nproc-per-node N scripts/film_reduce_check.py"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import pbrt_v4_b200 as P

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("gloo")  # only for the 128-byte id
torch.zeros(1, device="cuda")    # torch's CUDA context (and its NCCL, which the library then reuses)
import torch.cuda.nccl  # noqa: F401  (loads libnccl.so.2 into the process)
nccl = C.CDLL("libnccl.so.2", mode=C.RTLD_GLOBAL)
class UniqueId(C.Structure):  # ncclUniqueId: 128 bytes, passed BY VALUE to ncclCommInitRank
    _fields_ = [("internal", C.c_char * 128)]


uid = UniqueId()
if rank == 0:
    rc = nccl.ncclGetUniqueId(C.byref(uid))
    assert rc == 0, rc
box = [bytes(uid)]
dist.broadcast_object_list(box, src=0)
uid = UniqueId.from_buffer_copy(box[0])
comm = C.c_void_p()
nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
rc = nccl.ncclCommInitRank(C.byref(comm), world, uid, rank)
assert rc == 0, rc

rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", "cornell_spectral_r48_s16.b2rs"))
r = P.B200Renderer(rsc, device=torch.cuda.current_device())
first, n, stride = P.parallel.sample_stripe(rsc.spp, rank, world)
r.render(first, n, stride)
P.capi.check(P.capi.lib().b200pt_film_reduce(r._r, comm, 0, 0), r.agg._ctx)
r.agg.sync()
if rank == 0:
    got = [a.copy() for a in r.film_all() if a is not None]
    r.clear()
    r.render()
    want = [a for a in r.film_all() if a is not None]
    err = max(float(np.abs(g - w).max() / max(np.abs(w).max(), 1e-30)) for g, w in zip(got, want))
    print(f"film_reduce over {world} ranks: max relative difference to the single-GPU film {err:.3e}", flush=True)
    assert err <= 1e-12, err
    print("film_reduce ok", flush=True)
r.close()
dist.barrier()
nccl.ncclCommDestroy(comm)
dist.destroy_process_group()
