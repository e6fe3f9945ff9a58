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


def _simulate_peer_frames(world, depth, frames, seed, skip_ack_wait=False, split=True):
    """Discrete model of the peer-memory exchange: every GPU has `depth` in-order trace streams (+ one consumer stream when split), 128 flag words and `depth`
    framebuffers holding, per writer, the frame number last stored.  A random scheduler picks any stream whose head op can run.
    Returns None, or a string naming the first violated invariant."""
    import random
    rng = random.Random(seed)
    flags = [[0] * frame.PEER_WORDS for _ in range(world)]
    fb = [[[-1] * world for _ in range(depth)] for _ in range(world)]          # fb[gpu][slot][writer] = frame
    consumed = [[-1] * depth for _ in range(world)]                            # last frame consumed from fb[gpu][slot]
    streams = [dict() for _ in range(world)]                      # per GPU: stream name -> queue of ops
    recorded = [set() for _ in range(world)]
    for f in range(frames):
        for r in range(world):
            for op in frame.peer_frame_ops(f, depth, r, world, split=split):
                if skip_ack_wait and op[1] == "wait" and op[2][0] >= frame.PEER_ACK:
                    continue
                streams[r].setdefault(op[0], []).append(op[1:])
    pending = sum(len(q) for s in streams for q in s.values())
    while pending:
        ready = []
        for r in range(world):
            for name, q in streams[r].items():
                if not q:
                    continue
                op = q[0]
                if op[0] == "wait" and not all(flags[r][w] >= op[2] for w in op[1]):
                    continue
                if op[0] == "wait_event" and op[1] not in recorded[r]:
                    continue
                ready.append((r, name))
        if not ready:
            return "deadlock with %d ops pending" % pending
        r, name = rng.choice(ready)
        op = streams[r][name].pop(0)
        pending -= 1
        if op[0] == "store_pixels":
            _, slot, f = op
            for g in range(world):
                if f - depth >= 0 and consumed[g][slot] < f - depth:
                    return "GPU %d overwrote frame %d in slot %d of GPU %d before it was consumed" % (r, f - depth, slot, g)
                fb[g][slot][r] = f
        elif op[0] == "signal":
            for g in range(world):
                flags[g][op[1]] = op[2]
        elif op[0] == "record":
            recorded[r].add(op[1])
        elif op[0] == "consume":
            _, slot, f = op
            if fb[r][slot] != [f] * world:
                return "GPU %d consumed a torn frame %d: %r" % (r, f, fb[r][slot])
            consumed[r][slot] = f
    return None


@pytest.mark.parametrize("world,depth", [(2, 1), (2, 4), (3, 2), (4, 8), (8, 8), (8, 3)])
def test_peer_exchange_protocol_under_random_interleavings(world, depth):
    """The N > 1 path has no collective to run over gloo: its host-side logic IS the flag protocol (frame.peer_frame_ops = what
    clrt_render_frame_peers enqueues).  Random schedules of all GPUs' streams: never a deadlock, never a framebuffer overwritten
    before every GPU consumed it, never a torn frame.  Control: without the ack wait the model must catch an overwrite."""
    for split in (True, False):
        for seed in range(25):
            assert _simulate_peer_frames(world, depth, 5 * depth + 3, seed, split=split) is None
        caught = [_simulate_peer_frames(world, depth, 5 * depth + 3, seed, skip_ack_wait=True, split=split) for seed in range(25)]
        assert any(c is not None and "overwrote" in c for c in caught)


def test_peer_flag_words_do_not_collide():
    words = {frame.peer_word(kind, slot, w) for kind in (frame.PEER_READY, frame.PEER_ACK) for slot in range(16) for w in range(8)}
    assert len(words) == 256 and max(words) < frame.PEER_WORDS
    assert [frame.slot_sequence(f, 3) for f in range(7)] == [(0, 1), (1, 1), (2, 1), (0, 2), (1, 2), (2, 2), (0, 3)]
