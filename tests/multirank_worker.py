"""Worker of tests/test_multirank.py: launched by torch.distributed.run with the gloo backend.
Checks, on the CPU, the two ways the path shards over ranks:
  rays    - each rank traces its ray_shard with the host build of the device arithmetic
            (tests/native/hostcheck); the gathered answers equal the whole-batch answers bit for bit;
  samples - each rank renders its sample_stripe with the host loop over the render headers
            (tests/native/rendercheck); the gloo-reduced film equals the single-rank film."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pbrt_v4_b200 as P  # noqa: E402


def hostcheck(binary, scene_path, rays, mode):
    with tempfile.TemporaryDirectory() as tmp:
        rp, op = os.path.join(tmp, "r.bin"), os.path.join(tmp, "o.bin")
        rays.tofile(rp)
        subprocess.check_call([binary, scene_path, rp, mode, op], stdout=subprocess.DEVNULL)
        return np.fromfile(op, dtype=np.uint8)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    binary, rclib = sys.argv[1], sys.argv[2]

    # ---- rays
    sc = P.scenes.random_soup(3000, seed=5)
    rays = P.scenes.primary_rays(4001, seed=9)  # odd count: ragged shards
    with tempfile.TemporaryDirectory() as tmp:
        sp = os.path.join(tmp, "s.b2sc")
        sc.write(sp)
        a, b = P.parallel.ray_shard(len(rays), rank, world)
        mine = hostcheck(binary, sp, rays[a:b], "closest")
        whole = hostcheck(binary, sp, rays, "closest") if rank == 0 else None
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    if rank == 0:
        assert np.array_equal(np.concatenate(parts), whole), "sharded ray answers differ from the whole batch"

    # ---- samples
    import ctypes as C
    L = C.CDLL(rclib)
    rsc = P.RenderScene.load(os.path.join(ROOT, "tests", "golden", "cornell_r48_s8_uniform.b2rs"))
    w, h = rsc.resolution

    def run(first, n, stride):
        keep = []
        sd = P.render.scene_desc(rsc.geometry, None, keep)
        rd = rsc.desc(keep)
        film = np.zeros((h, w, 4), np.float64)
        msg = C.create_string_buffer(256)
        assert L.rendercheck_run(C.byref(sd), C.byref(rd), first, n, stride,
                                 film.ctypes.data_as(C.POINTER(C.c_double)), msg, 256, 2) == 0, msg.value
        return film

    stripes = [P.parallel.sample_stripe(rsc.spp, r, world) for r in range(world)]
    covered = sorted(f + k * s for f, n, s in stripes for k in range(n))
    assert covered == list(range(rsc.spp)), covered
    film = torch.from_numpy(run(*stripes[rank]))
    P.parallel.reduce_film(film, dist)
    if rank == 0:
        whole = run(0, rsc.spp, 1)
        assert np.array_equal(film.numpy()[..., 3], whole[..., 3])
        np.testing.assert_allclose(film.numpy(), whole, rtol=1e-12, atol=0)
        print("MULTIRANK_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
