"""Frame rendering across the GPUs of one box: the Demo's RTFrame() (R/src/Demos/OpenCLRayTracerDemo/
OpenCLRayTracerDemo.cpp:271-326) sharded in image space.  The reference is single-device; this is new.

One process per GPU (torch.distributed, NCCL).  The scene is static, so every rank builds the acceleration structure
once (redundant rebuild: a 1M-triangle LBVH build costs less than broadcasting its ~160 MB).  Per frame each rank
renders its stripe — 32x8-pixel tiles, tile t owned by rank t % world, persistent warps inside the GPU pull packets
from an atomic counter — into a device buffer in stripe order; ONE collective (all-gather of equal, padded RGBA8
stripes, 8.3 MB per 1080p frame) follows, and an untile kernel restores raster order.  No other exchange exists on
the path.  Eight frames are in flight: frames are traced on alternating streams (frame slots of the C ABI) so that the
latency-bound tail of one frame overlaps the next, and the collective, the untile and the read-back run on one more stream.

CLRT_EXCHANGE selects the exchange (BVH): "nccl" (the above; the default on one GPU), "fused" (the default at world > 1) or "push" —
both over peer memory (csrc/peer.cu): every rank exports its raster framebuffers + a block of 256 flag words (cudaIpc), and a frame's
pixels are stored straight into EVERY GPU's framebuffer over NVLink, by one pass over the stripe buffer ("push") or by the shading
kernel itself ("fused"); a "ready" word per (frame slot, rank) replaces the collective, an "ack" word orders the reuse of a slot's
framebuffer.  The waiting half of the exchange (everybody ready -> read-back -> ack) runs on its own stream (split exchange), so a
slot's trace stream never waits for the other GPUs' stripes.  On more than one GPU a stripe is ONE launch (frame_kernel: primary +
occlusion + shading; CLRT_FRAME_KERNELS), 16 frame slots; measured 8 B200s, Soup-1M 1080p: 0.325 ms per frame, 10.75 Grays/s, 7.05x
one GPU.

BounceRenderer (bottom of this file) is the configs[3] frame — primary + one diffuse bounce through the generic-ray entry point — with
three frames in flight on frame slots and an NCCL all-gather per frame.

PyTorch is plumbing here: device memory, the stream shared with NCCL, and the collective.
"""
import os
import sys

import numpy as np

from . import api, layouts as L

TILE_W, TILE_H, TILE_PIXELS = 32, 8, 256


def tile_grid(resX, resY):
    tx, ty = (resX + TILE_W - 1) // TILE_W, (resY + TILE_H - 1) // TILE_H
    return tx, ty, tx * ty


def stripe_capacity(resX, resY, world):
    """Slots (pixels) of one padded stripe: ceil(tiles / world) tiles of 256 slots."""
    _, _, tiles = tile_grid(resX, resY)
    return ((tiles + world - 1) // world) * TILE_PIXELS


def stripe_pixel_indices(resX, resY, rank, world):
    """Host mirror of the device stripe layout: for every slot of `rank`'s padded stripe the raster pixel index it
    holds, or -1 (outside the image / padding).  Slot = local_tile*256 + y_in_tile*32 + x_in_tile; local tile lt is
    global tile rank + lt*world."""
    tx, _, tiles = tile_grid(resX, resY)
    cap_tiles = (tiles + world - 1) // world
    lt = np.arange(cap_tiles)[:, None, None]
    gt = rank + lt * world
    x = (gt % tx) * TILE_W + np.arange(TILE_W)[None, None, :]
    y = (gt // tx) * TILE_H + np.arange(TILE_H)[None, :, None]
    pix = y * resX + x
    ok = (gt < tiles) & (x < resX) & (y < resY)
    return np.where(ok, pix, -1).reshape(-1).astype(np.int64)


def untile_host(stripes, resX, resY, world):
    """numpy mirror of clrt_frame_untile: gathered stripes (rank-major, each `stripe_capacity` slots) -> raster."""
    cap = stripe_capacity(resX, resY, world)
    stripes = np.asarray(stripes).reshape(world, cap)
    out = np.zeros(resX * resY, dtype=stripes.dtype)
    for r in range(world):
        idx = stripe_pixel_indices(resX, resY, r, world)
        ok = idx >= 0
        out[idx[ok]] = stripes[r][ok]
    return out.reshape(resY, resX)


def gather_stripes(local_stripe, world, group=None):
    """The one collective of the path: all-gather of equal padded stripes (torch tensor, any backend)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local_stripe
    out = torch.empty(world * local_stripe.numel(), dtype=local_stripe.dtype, device=local_stripe.device)
    dist.all_gather_into_tensor(out, local_stripe.contiguous(), group=group)
    return out


# ---- the peer-memory exchange protocol (csrc/peer.cu, clrt_render_frame_peers) in words -----------------------------------------
PEER_READY, PEER_ACK, PEER_WORDS = 0, 128, 256         # word offsets inside a GPU's flag block: [kind + slot*8 + writer rank]


def peer_word(kind, slot, writer):
    """Index of the flag word that GPU `writer` raises for frame slot `slot` (16 slots x 8 writers per kind)."""
    assert kind in (PEER_READY, PEER_ACK) and 0 <= slot < 16 and 0 <= writer < 8
    return kind + slot * 8 + writer


def slot_sequence(frame_no, depth):
    """Frame number -> (slot, sequence number carried by the slot's flag words; 1 for the slot's first frame)."""
    return frame_no % depth, frame_no // depth + 1


def peer_frame_ops(frame_no, depth, rank, world, split=True):
    """The stream-ordered steps one GPU enqueues for one frame, as (stream, op, ...) tuples — what clrt_render_frame_peers launches.
    stream = ("trace", slot) or ("consumer",).  split: the waiting half of the exchange (everybody's ready -> consume -> ack) runs on
    the consumer stream behind an event of this frame's own "ready"; otherwise everything is on the slot's trace stream.
    tests/test_frame_host.py runs these for all ranks under random interleavings: no deadlock, no framebuffer overwritten
    before every GPU consumed it, every consumed frame complete."""
    k, seq = slot_sequence(frame_no, depth)
    t, c = ("trace", k), (("consumer",) if split else ("trace", k))
    ops = []
    if seq > 1:
        ops.append((t, "wait", [peer_word(PEER_ACK, k, w) for w in range(world)], seq - 1))
    ops.append((t, "store_pixels", k, frame_no))                          # into every GPU's framebuffer k
    ops.append((t, "signal", peer_word(PEER_READY, k, rank), seq))       # on every GPU
    if split:
        ops.append((t, "record", frame_no))
        ops.append((c, "wait_event", frame_no))
    ops.append((c, "wait", [peer_word(PEER_READY, k, w) for w in range(world)], seq))
    ops.append((c, "consume", k, frame_no))                               # read-back / display
    ops.append((c, "signal", peer_word(PEER_ACK, k, rank), seq))
    return ops


class FrameRenderer:
    """Per-rank renderer: owns the context (on torch's current stream), the device scene and the built structure."""

    def __init__(self, scene_buf, resX, resY, accel="BVH", rank=0, world=1, device=0, group=None,
                 topLevelDensity=2.0, leafDensity=2.0):
        import torch
        self.torch = torch
        self.rank, self.world, self.group = rank, world, group
        self.resX, self.resY = resX, resY
        torch.cuda.set_device(device)
        self.device = torch.device("cuda", device)
        # one stream for everything of this renderer: our kernels (the context borrows it), torch's copies and NCCL's
        # ordering (torch orders the collective after/before work on the stream that is current when it is issued)
        self.stream = torch.cuda.Stream(self.device)
        assert self.stream.cuda_stream != 0
        self.ctx = api.Context(device, stream=self.stream.cuda_stream)
        self.scene = api.Scene(scene_buf)
        self.scene.loadToGPU(self.ctx)
        if accel.upper() == "BVH":                     # -accStruct {BVH,GRID} (OpenCLRayTracerDemo.cpp:66-73); KD is ours
            self.accel = api.BVHManager(self.ctx, self.scene)
        elif accel.upper() == "KD":
            self.accel = api.KDTreeManager(self.ctx, self.scene)
        else:
            self.accel = api.TwoLevelGridManager(self.ctx, self.scene, topLevelDensity, leafDensity)
        self.accel.initialize()
        if accel.upper() == "GRID" and os.environ.get("CLRT_GRID_WALK", "corrected") == "corrected":
            self.accel.set_trace_mode(api.TRACE_FAST)  # the corrected closest-hit walk (the literal one is the grid's default)
        # Kernels per frame.  Two (primary contacts, then shadow + shade) are faster on a whole frame and on half a frame (2.31 against
        # 2.45 ms on one GPU: warps that mix closest-hit and occlusion rays lose coherence).  ONE kernel for the whole stripe
        # (clrt_accel_set_fused_shadows: 2) wins on small stripes: measured on 8 B200s, Soup-1M, 0.330 ms per frame against 0.399 with two
        # kernels, 0.631 against 0.743 on 4 and 1.249 against 1.446 on 2 (profiles/r02_bench_soup1m_*gpu.json; the single-GPU stripe
        # probe agrees for one kernel, profiles/r02_stripe_probe.md) — so it is the default on more than one GPU.
        # CLRT_FRAME_KERNELS=1|2 overrides.
        fk = os.environ.get("CLRT_FRAME_KERNELS", "auto")
        self.one_kernel = accel.upper() == "BVH" and (fk == "1" or (fk == "auto" and world >= 2))
        if self.one_kernel:
            self.accel.set_fused_shadows(2)
        self.cap = stripe_capacity(resX, resY, world)
        torch.cuda.synchronize()
        # DEPTH frames in flight.  Frame k is traced on trace stream k % DEPTH with frame slot k % DEPTH of the C ABI (private
        # work counter + scratch), so the tail of one frame's persistent kernels — a few long rays on a mostly idle GPU —
        # overlaps the head of the next frame's; gather + untile (+ read-back) run on one more stream (`comm`), in frame
        # order on every rank (the collective's order must agree across ranks).  Buffers are per slot; events order reuse.
        self.split_exchange = os.environ.get("CLRT_SPLIT_EXCHANGE", "1") != "0"
        # Frames per pair of launches with the fused exchange (clrt_render_frames_peers; two-kernel mode): measured on 8 GPUs it
        # bought 1 % (0.394 against 0.399 ms) and cost the e2e path its overlap, so it is off unless CLRT_FRAME_BATCH=2..4 asks for it.
        fb = os.environ.get("CLRT_FRAME_BATCH", "1")
        self.batch = 1 if fb == "auto" else max(1, min(4, int(fb)))
        if accel.upper() != "BVH" or not self.split_exchange or self.one_kernel:
            self.batch = 1
        depth = int(os.environ.get("CLRT_FRAME_DEPTH", str(8 if self.batch == 1 else 4 * self.batch)))      # four launches' worth of frames in flight
        self.DEPTH = max(1, min(16, depth)) if accel.upper() == "BVH" else 1
        if self.DEPTH % self.batch:
            self.batch = 1
        self.comm = torch.cuda.Stream(self.device)
        self.trace_streams = [self.stream] + [torch.cuda.Stream(self.device) for _ in range(self.DEPTH - 1)]
        self.stripes = [torch.zeros(self.cap, dtype=torch.int32, device=self.device) for _ in range(self.DEPTH)]
        self.gathereds = [torch.zeros(self.cap * world, dtype=torch.int32, device=self.device) if world > 1 else self.stripes[i]
                          for i in range(self.DEPTH)]
        self.rasters = [torch.zeros(resY * resX, dtype=torch.int32, device=self.device) for _ in range(self.DEPTH)]
        self.host_rasters = [torch.zeros(resY * resX, dtype=torch.int32).pin_memory() for _ in range(self.DEPTH)]
        self.ev_traced = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.ev_free = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.ev_host = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.frame_no = 0
        self.exchange = os.environ.get("CLRT_EXCHANGE", "fused" if world > 1 else "nccl").lower() if accel.upper() == "BVH" else "nccl"
        if self.exchange not in ("nccl", "push", "fused"):
            raise ValueError("CLRT_EXCHANGE=%r (nccl, push or fused)" % self.exchange)
        self._peer_block = None
        self._slot, self._slot_seq, self._pending = 0, [0] * 16, []
        if self.exchange != "nccl":
            self._setup_peers()
        self.stripe, self.gathered, self.raster, self.host_raster = self.stripes[0], self.gathereds[0], self.rasters[0], self.host_rasters[0]

    def build(self):
        self.join()
        self.accel.initializeFrame()
        self.accel.construct()
        self.stream.synchronize()

    # ---- exchange over peer memory ------------------------------------------------------------------------------------------

    def _setup_peers(self):
        """One exportable block per rank: DEPTH raster framebuffers + 1 KiB of flag words.  Handles travel by all_gather_object.
        Every rank must end up in the same mode: if any rank cannot export or open (no peer access between two GPUs), all of
        them fall back to the NCCL exchange, and say so."""
        import torch.distributed as dist
        torch = self.torch
        raster_bytes = (self.resX * self.resY * 4 + 255) & ~255

        def everyone(ok):
            if self.world == 1:
                return ok
            oks = [None] * self.world
            dist.all_gather_object(oks, bool(ok), group=self.group)
            return all(oks)

        handle, why = None, ""
        try:
            self._peer_block = api.peer_alloc(self.ctx, self.DEPTH * raster_bytes + 1024)
            handle = api.peer_export(self.ctx, self._peer_block)
        except api.ClrtError as e:
            why = str(e)
        handles = [handle]
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, handle, group=self.group)
        self._peer_bases = [None] * self.world
        if all(h is not None for h in handles):
            try:
                for r in range(self.world):
                    self._peer_bases[r] = self._peer_block if r == self.rank else api.peer_open(self.ctx, handles[r])
            except api.ClrtError as e:
                why = str(e)
        if not everyone(all(b is not None for b in self._peer_bases)):
            for r, b in enumerate(self._peer_bases):
                if b is not None and r != self.rank:
                    api.peer_close(self.ctx, b)
            if self._peer_block is not None:
                api.peer_free(self.ctx, self._peer_block)
            self._peer_block = None
            print("[clrt] rank %d: peer-memory exchange unavailable (%s); using the NCCL exchange" % (self.rank, why or "a peer failed"),
                  file=sys.stderr, flush=True)
            self.exchange = "nccl"
            return
        self._peer_rasters = [api.peer_pointer_array([b + k * raster_bytes for b in self._peer_bases]) for k in range(self.DEPTH)]
        self._peer_flags = api.peer_pointer_array([b + self.DEPTH * raster_bytes for b in self._peer_bases])
        self._own_flags = self._peer_block + self.DEPTH * raster_bytes

        class _Raw:                                       # torch view of our own allocation (for the read-back and the tests)
            def __init__(self, ptr, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 2}
        self.rasters = [torch.as_tensor(_Raw(self._peer_block + k * raster_bytes, self.resX * self.resY), device=self.device)
                        for k in range(self.DEPTH)]
        if self.world > 1:
            dist.barrier(group=self.group)                # nobody writes before everybody has opened

    def _take_slots(self, m):
        """m consecutive frame slots starting at the current one and their sequence numbers (per-slot counters: every rank makes the
        same calls in the same order, so they agree across the box)."""
        k0 = self._slot
        assert k0 + m <= self.DEPTH
        self._slot = (k0 + m) % self.DEPTH
        seqs = []
        for k in range(k0, k0 + m):
            self._slot_seq[k] += 1
            seqs.append(self._slot_seq[k])
        self.frame_no += m
        return k0, seqs

    def _flush(self):
        """Submits the pending frames (1..batch, consecutive slots) as ONE pair of launches (clrt_render_frames_peers)."""
        pend, self._pending = self._pending, []
        if not pend:
            return
        k0, seqs = self._take_slots(len(pend))
        for i, (_, to_host) in enumerate(pend):
            if to_host:
                self.ev_host[k0 + i].synchronize()
        ts = self.trace_streams[k0]
        hosts = [self.host_rasters[k0 + i].data_ptr() if (p[1] and self.rank == 0) else None for i, p in enumerate(pend)]
        self.accel.render_frames_peers(ts.cuda_stream, k0, [p[0] for p in pend], self.rank, self.world,
                                       [self._peer_rasters[k0 + i] for i in range(len(pend))], self._peer_flags, seqs, hosts, self.comm.cuda_stream)
        for i, p in enumerate(pend):
            if p[1]:
                self.ev_host[k0 + i].record(self.comm)

    def _submit_peers(self, camera, shadows, want_rays, to_host):
        torch = self.torch
        if self.batch > 1 and self.exchange == "fused" and shadows and not want_rays:
            # frames are collected and traced `batch` at a time (one pair of launches for all their stripes); join() / finish() flush
            if not self._pending and self._slot + self.batch > self.DEPTH:
                self._slot = 0                              # a batch occupies consecutive slots
            k = self._slot + len(self._pending)
            self._pending.append((camera, to_host))
            self.stripe, self.raster, self.host_raster = self.stripes[k], self.rasters[k], self.host_rasters[k]
            if len(self._pending) == self.batch:
                self._flush()
            return None
        self._flush()
        k, seqs = self._take_slots(1)
        seq = seqs[0]
        if to_host:
            self.ev_host[k].synchronize()
        ts = self.trace_streams[k]
        s = ts.cuda_stream
        if self.exchange == "fused" and shadows:          # the whole protocol below in one C call
            host = self.host_rasters[k].data_ptr() if (to_host and self.rank == 0) else None
            # the exchange's waiting half (everybody's ready -> read-back -> ack) runs on the `comm` stream: the slot's trace stream
            # never waits for the other GPUs' stripes, nor for rank 0's 8 MB read-back
            rays = self.accel.render_frame_peers(s, k, camera, self.rank, self.world, self._peer_rasters[k], self._peer_flags, seq, host, want_rays,
                                                 consumer_stream=self.comm.cuda_stream if self.split_exchange else None)
            if to_host:
                self.ev_host[k].record(self.comm if self.split_exchange else ts)
            self.stripe, self.raster, self.host_raster = self.stripes[k], self.rasters[k], self.host_rasters[k]
            return rays
        with torch.cuda.stream(ts):
            if seq > 1:                                   # every GPU has consumed the previous frame of slot k from its framebuffer
                api.peer_wait(self.ctx, s, self._own_flags, peer_word(PEER_ACK, k, 0), self.world, seq - 1)
            rays = self.accel.render_stripe_on(s, k, camera, shadows, self.rank, self.world, self.stripes[k].data_ptr(), want_rays)
            api.frame_push(self.ctx, s, self.stripes[k].data_ptr(), self.resX, self.resY, self.rank, self.world, self._peer_rasters[k])
            api.peer_signal(self.ctx, s, self._peer_flags, self.world, peer_word(PEER_READY, k, self.rank), seq)
            api.peer_wait(self.ctx, s, self._own_flags, peer_word(PEER_READY, k, 0), self.world, seq)
            if to_host and self.rank == 0:
                self.host_rasters[k].copy_(self.rasters[k], non_blocking=True)
            api.peer_signal(self.ctx, s, self._peer_flags, self.world, peer_word(PEER_ACK, k, self.rank), seq)
            if to_host:
                self.ev_host[k].record(ts)
        self.stripe, self.raster, self.host_raster = self.stripes[k], self.rasters[k], self.host_rasters[k]
        return rays

    def _submit(self, camera, shadows, want_rays, to_host):
        if self.exchange != "nccl":
            return self._submit_peers(camera, shadows, want_rays, to_host)
        torch = self.torch
        k = self.frame_no % self.DEPTH
        self.frame_no += 1
        if to_host:
            self.ev_host[k].synchronize()             # the host buffer of slot k may still be the target of frame k-DEPTH's copy
        ts = self.trace_streams[k]
        with torch.cuda.stream(ts):
            ts.wait_event(self.ev_free[k])            # slot k's stripe buffer is free once frame k-DEPTH was gathered
            if self.DEPTH > 1:
                rays = self.accel.render_stripe_on(ts.cuda_stream, k, camera, shadows, self.rank, self.world, self.stripes[k].data_ptr(), want_rays)
            else:
                rays = self.accel.render_stripe(camera, shadows, self.rank, self.world, self.stripes[k].data_ptr(), want_rays)
            self.ev_traced[k].record(ts)
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self.ev_traced[k])
            if self.world > 1:
                import torch.distributed as dist
                dist.all_gather_into_tensor(self.gathereds[k], self.stripes[k], group=self.group)
            api.frame_untile_on(self.ctx, self.comm.cuda_stream, self.gathereds[k].data_ptr(), self.resX, self.resY, self.world,
                                self.rasters[k].data_ptr())
            if to_host and self.rank == 0:
                self.host_rasters[k].copy_(self.rasters[k], non_blocking=True)
            self.ev_free[k].record(self.comm)
            if to_host:
                self.ev_host[k].record(self.comm)
        self.stripe, self.gathered, self.raster, self.host_raster = self.stripes[k], self.gathereds[k], self.rasters[k], self.host_rasters[k]
        return rays

    def render_device(self, camera, shadows=True, want_rays=False):
        """Submits one frame; its raster-ordered result (every rank holds the whole frame) is in `self.raster` once the
        second stream has passed — `finish()` joins — and stays there until frame k+DEPTH is submitted (by any rank, with the
        peer-memory exchanges).  Returns the rays traced by this rank if want_rays (a host sync)."""
        return self._submit(camera, shadows, want_rays, False)

    def render_host_async(self, camera, shadows=True):
        """e2e, pipelined: host camera in, RGBA8 raster into pinned host memory (rank 0); returns the host buffer the
        frame will land in — valid after `finish()` or once two more frames were submitted."""
        self._submit(camera, shadows, False, True)
        return self.host_raster

    def render_host(self, camera, shadows=True):
        """e2e, synchronous: returns when the frame is in pinned host memory (rank 0 reads the frame back)."""
        out = self.render_host_async(camera, shadows)
        self.finish()
        return out

    def join(self):
        """Makes the context's stream wait for every frame submitted so far (device-side join, no host wait)."""
        if self.exchange != "nccl":
            self._flush()                               # a batch that did not fill up
        for ts in self.trace_streams[1:]:
            self.stream.wait_stream(ts)
        self.stream.wait_stream(self.comm)

    def finish(self):
        """Joins all streams into the context's stream and waits for it."""
        self.join()
        self.stream.synchronize()

    def close(self):
        if self._peer_block is not None:
            import torch.distributed as dist
            self.finish()
            if self.world > 1:
                dist.barrier(group=self.group)            # no peer still writes into (or reads flags from) our block
            for r, b in enumerate(self._peer_bases):
                if r != self.rank:
                    api.peer_close(self.ctx, b)
            self.rasters = []
            api.peer_free(self.ctx, self._peer_block)
            self._peer_block = None
        self.accel.free(); self.scene.free(); self.ctx.close()


def orbit_camera(scene_buf, resX, resY, step, steps_per_turn=64, fov_deg=90.0):
    """Demo camera (OpenCLRayTracerDemo.cpp:202-217) moved along a small arc so successive frames differ:
    yaw about the scene centre by +-0.2 rad, same distance.  float32 like the reference's Transform.h."""
    hdr = np.frombuffer(scene_buf[:96].tobytes(), dtype=L.SCENE_HEADER)[0]
    bb = hdr["modelsBoundingBox"].astype(np.float64)
    centre = (bb[0, :3] + bb[1, :3]) * 0.5
    half = (bb[1, :3] - bb[0, :3]) * 0.5
    base = api.demo_camera(scene_buf, resX, resY, fov_deg)
    ang = 0.2 * np.sin(2.0 * np.pi * step / steps_per_turn)
    dist = 2.0 * half[2]
    pos = np.array([centre[0] + dist * np.sin(ang), centre[1], centre[2] - dist * np.cos(ang)])
    # camera looks along +z in its own frame; yaw so that it faces the centre.  Rows of the 3x4 matrix (Transform.h).
    c, s = np.cos(-ang), np.sin(-ang)
    m = np.array([c, 0, s, pos[0], 0, 1, 0, pos[1], -s, 0, c, pos[2]], dtype=np.float32)
    return api.make_camera(resX, resY, base["FOVDistance"], m)


class BounceRenderer:
    """BASELINE.json configs[3]: incoherent secondary rays, image-tile sharded.  Per frame and rank: primary contacts of the rank's
    stripe -> one cosine-weighted diffuse bounce ray per hit -> the generic-ray entry generateContacts(rays, contacts, n)
    (BVHManager.cpp:292-329) -> Blinn-Phong halved where the bounce ray hit -> RGBA8 stripe (clrt_render_bounce_stripe_on = the four
    public calls on a frame slot); one all-gather of equal padded stripes + untile gives every rank the frame.  Rays are independent and
    the structure is read-only: no other exchange exists.  DEPTH frames are in flight (CLRT_BOUNCE_DEPTH, default 3): every frame slot has
    its own stream, work counter and buffers, so the all-gather and the kernel tails of one frame overlap the tracing of the next; the
    collectives are issued in frame order on every rank."""

    def __init__(self, scene_buf, resX, resY, rank=0, world=1, device=0, group=None):
        import torch
        self.torch = torch
        self.rank, self.world, self.group = rank, world, group
        self.resX, self.resY = resX, resY
        torch.cuda.set_device(device)
        self.device = torch.device("cuda", device)
        self.stream = torch.cuda.Stream(self.device)
        self.ctx = api.Context(device, stream=self.stream.cuda_stream)
        self.scene = api.Scene(scene_buf)
        self.scene.loadToGPU(self.ctx)
        self.accel = api.BVHManager(self.ctx, self.scene)
        self.accel.initialize()
        self.cap, self.tiles = api.frame_stripe_pixels(resX, resY, rank, world)
        self.slots = self.tiles * TILE_PIXELS
        self.DEPTH = max(1, min(8, int(os.environ.get("CLRT_BOUNCE_DEPTH", "3"))))
        self.frames = []
        for k in range(self.DEPTH):
            st = torch.cuda.Stream(self.device)
            with torch.cuda.stream(st):
                f = {"stream": st, "event": torch.cuda.Event(),
                     "contacts": torch.zeros(self.cap * 32, dtype=torch.uint8, device=self.device),
                     "rays": torch.zeros(self.cap * 48, dtype=torch.uint8, device=self.device),
                     "bounce": torch.zeros(self.cap * 32, dtype=torch.uint8, device=self.device),
                     "stripe": torch.zeros(self.cap, dtype=torch.int32, device=self.device),
                     "raster": torch.zeros(resY * resX, dtype=torch.int32, device=self.device),
                     "host_raster": torch.zeros(resY * resX, dtype=torch.int32).pin_memory() if rank == 0 else None}
                f["gathered"] = torch.zeros(self.cap * world, dtype=torch.int32, device=self.device) if world > 1 else f["stripe"]
            st.synchronize()
            self.frames.append(f)
        self._next, self._last = 0, 0
        self.stream.synchronize()

    # the buffers of the frame submitted last (bench.py's per-kernel timing and ray count read them)
    contacts = property(lambda self: self.frames[self._last]["contacts"])
    rays = property(lambda self: self.frames[self._last]["rays"])
    bounce = property(lambda self: self.frames[self._last]["bounce"])
    raster = property(lambda self: self.frames[self._last]["raster"])
    host_raster = property(lambda self: self.frames[self._last]["host_raster"])

    def build(self):
        self.finish()
        self.accel.initializeFrame()
        self.accel.construct()
        self.stream.synchronize()

    def render_device(self, camera, seed=20161116):
        torch = self.torch
        k = self._next
        self._next, self._last = (k + 1) % self.DEPTH, k
        f = self.frames[k]
        with torch.cuda.stream(f["stream"]):
            if self.slots:
                self.accel.render_bounce_stripe_on(f["stream"].cuda_stream, k, camera, self.rank, self.world, seed, f["contacts"].data_ptr(),
                                                   f["rays"].data_ptr(), f["bounce"].data_ptr(), f["stripe"].data_ptr())
            if self.world > 1:
                import torch.distributed as dist
                dist.all_gather_into_tensor(f["gathered"], f["stripe"], group=self.group)
            api.frame_untile_on(self.ctx, f["stream"].cuda_stream, f["gathered"].data_ptr(), self.resX, self.resY, self.world, f["raster"].data_ptr())
            f["event"].record(f["stream"])
        return f

    def render_host(self, camera, seed=20161116):
        """e2e: host camera in, the frame in pinned host memory (rank 0) when the frame's stream has passed."""
        f = self.render_device(camera, seed)
        if self.rank == 0:
            with self.torch.cuda.stream(f["stream"]):
                f["host_raster"].copy_(f["raster"], non_blocking=True)
                f["event"].record(f["stream"])
        return f["host_raster"]

    def join(self):
        """Device-side join: the renderer's own stream waits for every frame in flight."""
        for f in self.frames:
            self.stream.wait_event(f["event"])

    def count_rays(self):
        """Rays of the last frame on this rank: primary = stripe slots inside the image, bounce = primary hits (host sync)."""
        self.finish()
        c = self.contacts[: self.slots * 32].view(self.torch.int32).reshape(-1, 8)
        inside = int((c[:, 0] != -1).sum().item())
        hits = int((c[:, 7].view(self.torch.float32) > 0).sum().item())
        return inside + hits

    def finish(self):
        for f in self.frames:
            f["stream"].synchronize()
        self.stream.synchronize()

    def close(self):
        self.finish()
        self.accel.free(); self.scene.free(); self.ctx.close()
