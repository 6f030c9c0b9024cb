"""Multi-GPU rendering: one process per GPU, samples split by index, ONE reduce of the float accumulation buffer.

The reference is a single process (rayon over pixels, renderer/mod.rs:105-106) whose samples are independent
because its RNG is seeded per (pixel, sample) (mod.rs:110-112).  Here rank g of G renders the samples
{s : s mod G == g} of the FULL frame into its own W*H*3 f32 buffer, the buffers are summed onto rank 0 with a
single `torch.distributed.reduce` (NCCL over NVLink on GPUs, gloo in the CPU tests), and rank 0 runs
mean -> tonemap -> gamma -> u8 (post_processing, mod.rs:335-353).  The union of samples is identical for any G, so
the image differs between G only by the f32 summation order.

No collective is issued inside a frame; the reduce is the only exchange step of the path.
"""
import os

import torch
import torch.distributed as dist


def sample_partition(total_samples, world_size, rank):
    """Samples of `rank`: (begin, stride, count) with sample indices begin + k*stride, k < count."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    count = (total_samples - rank + world_size - 1) // world_size if total_samples > rank else 0
    return rank, world_size, count


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def reduce_accum(accum, dst=0):
    """Sum the per-rank accumulation buffers onto `dst` (in place there); a no-op for a single process."""
    rank, size = world()
    if size > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def broadcast_arrays(tensors, src=0, group=None):
    """Send every tensor of `tensors` from rank `src` to the same-sized tensor of the other ranks (in place)."""
    rank, size = world()
    if size > 1:
        for t in tensors:
            dist.broadcast(t, src=src, group=group)
    return tensors


def part_size(nbytes, parts):
    """Bytes of one part of pt_scene_upload_part: (size / parts) rounded down to a multiple of 256."""
    return (nbytes // parts) & ~255


def allgather_parts(tensors, group=None):
    """In-place all-gather of byte tensors whose part r (part_size bytes at offset r * part_size) is valid on rank r;
    the remainder behind the last part is not touched (every rank has it already)."""
    rank, size = world()
    if size > 1:
        for t in tensors:
            n = part_size(t.numel(), size)
            if n:
                dist.all_gather_into_tensor(t[:n * size], t[rank * n:(rank + 1) * n], group=group)
    return tensors


class _DeviceBytes:
    """A raw device allocation as a __cuda_array_interface__ object (torch.as_tensor wraps it without copying)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def render_distributed(render_fn, tonemap_fn, total_samples, accum):
    """Generic driver used by both the CUDA path and the CPU (gloo) tests.

    render_fn(sample_begin, sample_stride, n_samples, accum) adds this rank's samples into `accum`;
    tonemap_fn(accum, total_samples) -> image runs on rank 0 only.  Returns the image on rank 0, None elsewhere.
    """
    rank, size = world()
    begin, stride, count = sample_partition(total_samples, size, rank)
    accum.zero_()
    render_fn(begin, stride, count, accum)
    reduce_accum(accum, 0)
    if rank == 0:
        return tonemap_fn(accum, total_samples)
    return None


class DistributedRenderer:
    """Renderer::render (renderer/mod.rs:76-169) across the ranks of the default process group.

    Every rank holds its own replica of the scene on its own GPU (`scene` was created with device=LOCAL_RANK).
    """

    def __init__(self, scene, profile, seed=0, device=None):
        from . import api
        self.api = api
        self.scene, self.profile, self.seed = scene, profile, seed
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        h, w = profile.height, profile.width
        self.accum = torch.zeros(h, w, 3, dtype=torch.float32, device=self.device)
        self.rgb = torch.zeros(h, w, 3, dtype=torch.uint8, device=self.device)
        self.last_stats = None
        self._dist_tensors = {}

    def distribute_scene(self, mode=None, src=0):
        """Scene::load's result reaching every GPU, once per frame.  Returns the bytes THIS rank read from its host.

        mode "replicate": every rank re-sends the whole scene from its own pinned host copy (pt_scene_upload) - N
            uploads of the same bytes; fine while N x bytes per frame stays below what one host's PCIe / memory
            system delivers (2 GPUs: e2e within 1.2 % of the device-resident figure).
        mode "allgather": rank r sends part r of N of every array (pt_scene_upload_part) and one in-place NCCL
            all-gather per array replicates it over NVLink / NVSwitch: the scene crosses the host's PCIe once per
            frame.  Eight uploads of the console scene are 121 GB/s against the 107 GB/s the 8-GPU box delivered
            (profiles/r07/s26_8gpu_bench.json: 11.1 ms per step against a 9.7 ms frame).
        mode "broadcast": rank `src` uploads, one NCCL broadcast per array.
        Default ($PT_B200_SCENE_DIST): "allgather" from 8 ranks up, "replicate" below.  The host-to-device copies run
        on the scene's copy stream (they overlap the frame the GPU is still rendering); the collectives are enqueued on
        the current stream, between two frames - beside a frame they wait for SM slots that the persistent kernels
        hold (measured at 2 GPUs: +2.1 ms per step on a stream of their own, +1.4 ms in line)."""
        rank, size = world()
        if mode is None:
            mode = os.environ.get("PT_B200_SCENE_DIST") or ("allgather" if size >= 8 else "replicate")
        if size == 1 or mode == "replicate":
            return self.scene.upload()
        assert mode in ("allgather", "broadcast"), mode
        nbytes = 0
        if mode == "allgather":
            nbytes = self.scene.upload_part(rank, size)
        elif rank == src:
            nbytes = self.scene.upload()
        raw = torch.cuda.current_stream(self.device).cuda_stream
        targets = self.scene.upload_targets(raw)
        key = tuple(targets)
        if key not in self._dist_tensors:
            self._dist_tensors[key] = [torch.as_tensor(_DeviceBytes(p, n), device=self.device) for p, n in targets if n]
        if mode == "allgather":
            allgather_parts(self._dist_tensors[key])
        else:
            broadcast_arrays(self._dist_tensors[key], src)
        self.scene.upload_commit(raw)
        return nbytes

    def _render(self, begin, stride, count, accum, want_stats=False):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        self.last_stats = self.api.render_accum_device(self.scene, self.profile, accum.data_ptr(), begin, stride, count,
                                                       self.seed, stream, want_stats)

    def _tonemap(self, accum, total_samples):
        stream = torch.cuda.current_stream(self.device).cuda_stream
        p = self.profile
        self.api.tonemap_device(accum.data_ptr(), p.width, p.height, total_samples, p.tonemap, self.rgb.data_ptr(),
                                self.device.index or 0, stream)
        return self.rgb

    def render_phases(self, total_samples=None):
        """One instrumented frame: CUDA-event milliseconds of this rank's render, of the reduce as this rank sees it
        (it includes waiting for the slowest rank) and of the tonemap (rank 0).  Synchronises; fills last_stats."""
        total = self.profile.samples if total_samples is None else total_samples
        s = torch.cuda.current_stream(self.device)
        rank, size = world()
        begin, stride, count = sample_partition(total, size, rank)
        out = None
        for _ in range(2):  # the first pass brings the device back from the idle state a barrier leaves it in
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            self.accum.zero_()
            ev[0].record(s)
            self._render(begin, stride, count, self.accum, want_stats=True)
            ev[1].record(s)
            reduce_accum(self.accum, 0)
            ev[2].record(s)
            if rank == 0:
                self._tonemap(self.accum, total)
            ev[3].record(s)
            s.synchronize()
            out = {"ms_render": ev[0].elapsed_time(ev[1]), "ms_reduce": ev[1].elapsed_time(ev[2]),
                   "ms_tonemap": ev[2].elapsed_time(ev[3]), "ms_render_device": self.last_stats["ms_total"]}
        return out

    def render(self, total_samples=None, want_stats=False):
        """Asynchronous on the current stream; returns the device RGB8 image on rank 0 (None elsewhere)."""
        total = self.profile.samples if total_samples is None else total_samples
        return render_distributed(lambda b, s, c, a: self._render(b, s, c, a, want_stats), self._tonemap, total, self.accum)
