"""Host logic of the image-space sharding (openclraytracer_b200/frame.py), no GPU: stripe layout vs the C ABI's own
bookkeeping, numpy untile, and the N > 1 path (one collective: all-gather of padded stripes) over gloo with world_size 2."""
import os
import sys
import numpy as np
import pytest
from conftest import ROOT
from openclraytracer_b200 import frame


@pytest.mark.parametrize("res", [(1920, 1080), (640, 480), (97, 61), (32, 8), (5, 3)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_stripes_partition_the_image(res, world):
    from openclraytracer_b200 import api
    seen = np.zeros(res[0] * res[1], dtype=np.int32)
    cap = frame.stripe_capacity(res[0], res[1], world)
    for r in range(world):
        idx = frame.stripe_pixel_indices(res[0], res[1], r, world)
        assert len(idx) == cap
        c_cap, c_tiles = api.frame_stripe_pixels(res[0], res[1], r, world)         # host-only entry point of the C ABI
        assert c_cap == cap
        assert c_tiles == len(np.unique(np.nonzero(idx.reshape(-1, 256) >= 0)[0])) or res[0] * res[1] < 256
        np.add.at(seen, idx[idx >= 0], 1)
    assert (seen == 1).all()                                                       # every pixel owned exactly once


def test_untile_host_inverts_the_stripe_layout():
    resX, resY, world = 100, 37, 3
    stripes = []
    for r in range(world):
        idx = frame.stripe_pixel_indices(resX, resY, r, world)
        stripes.append(np.where(idx >= 0, idx, 0xFFFFFF).astype(np.int32))
    img = frame.untile_host(np.concatenate(stripes), resX, resY, world)
    assert np.array_equal(img.reshape(-1), np.arange(resX * resY))


def _worker(rank, world, port, resX, resY, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from openclraytracer_b200 import frame as F
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    idx = F.stripe_pixel_indices(resX, resY, rank, world)
    local = torch.from_numpy(np.where(idx >= 0, idx * 7 + 1, 0).astype(np.int32))     # a fake "render" of this stripe
    gathered = F.gather_stripes(local, world)
    img = F.untile_host(gathered.numpy(), resX, resY, world)
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), img)
    dist.destroy_process_group()


def test_two_rank_gather_over_gloo(tmp_path):
    import torch.multiprocessing as mp
    resX, resY, world = 130, 70, 2
    port = 29500 + (os.getpid() % 500)
    mp.spawn(_worker, args=(world, port, resX, resY, str(tmp_path)), nprocs=world, join=True)
    want = (np.arange(resX * resY) * 7 + 1).reshape(resY, resX)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"rank{r}.npy"), want)           # every rank ends with the whole frame


def test_orbit_camera_step0_is_the_demo_camera():
    from openclraytracer_b200 import api, scenes
    buf = scenes.load_bundled("cube")
    a, b = frame.orbit_camera(buf, 64, 48, 0), api.demo_camera(buf, 64, 48)
    assert np.allclose(a["viewTransform"], b["viewTransform"], atol=1e-6) and a["FOVDistance"] == b["FOVDistance"]
