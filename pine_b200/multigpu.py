"""Multi-GPU sharding of one `PathIntegrator::render`: one process per GPU, the scene
replicated, (pixel, sample) work split, ONE reduce(sum) of the float film over NCCL at the end.

Every (pixel, sample) of path.cpp:31-40 is independent and every pixel has one writer, so there is
no exchange step inside the render; the only collective is the final film reduction.

  * mode "samples" (BASELINE.json's C5): rank r renders samples [r*spp/G, (r+1)*spp/G) of every
    pixel; its film holds sum/spp over its samples and alpha = n/spp, so the rank films ADD to the
    full film. Per-pixel RNG streams are re-seeded per rank with hash(p, first_sample) -- the
    reference's own `start_pixel(p, sample_index)` argument (sampler.h:286-289); fp32 sum order
    differs from one GPU, so 1-GPU and G-GPU films agree to rounding, not bit for bit.
  * mode "rows": the film is cut into bands of `band` rows dealt round-robin to the ranks (fine
    interleave against the light-facing-region imbalance); every rank renders all samples of its
    bands, other pixels stay 0, so the same reduce(sum) assembles a film that is BIT-IDENTICAL to
    the single-GPU render.

The plan functions are pure (tested on CPU); `ShardedRenderer` drives a `Backend` and reduces with
`torch.distributed` (NCCL on GPUs; the gloo tests substitute a CPU film model for the backend).
"""
from . import _abi as A

PB2_RENDER_KEEP_FILM = 4


def sample_shards(spp, world):
    """[(begin, end)] per rank; contiguous, covering [0, spp), sizes differ by at most one."""
    if spp <= 0 or world <= 0:
        raise ValueError("spp and world must be positive")
    base, extra = divmod(spp, world)
    out, b = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((b, b + n))
        b += n
    return out


def row_bands(height, world, rank, band=16):
    """[(row_begin, row_end)] of the bands dealt to `rank`: band k goes to rank k % world."""
    if band <= 0:
        raise ValueError("band must be positive")
    n_bands = (height + band - 1) // band
    return [(k * band, min(height, (k + 1) * band)) for k in range(rank, n_bands, world)]


class ShardedRenderer:
    """Renders this rank's share on its Backend and reduces the film to rank `dst`."""

    def __init__(self, backend, rank, world, mode="samples", band=16, group=None):
        if mode not in ("samples", "rows"):
            raise ValueError("mode must be 'samples' or 'rows'")
        self.backend, self.rank, self.world, self.mode, self.band, self.group = backend, rank, world, mode, band, group

    def plan(self, spp):
        """[(rows, samples, flags)] render calls of this rank."""
        w, h = self.backend.size
        if self.mode == "samples":
            s = sample_shards(spp, self.world)[self.rank]
            return [((0, h), s, 0)] if s[1] > s[0] else []
        bands = row_bands(h, self.world, self.rank, self.band)
        return [(b, (0, spp), PB2_RENDER_KEEP_FILM if i else 0) for i, b in enumerate(bands)]

    def _torch_stream(self, stream):
        """The ONE stream the render, the staging copies and the NCCL collectives of this rank are ordered on.
        The context's own stream is created cudaStreamNonBlocking and NCCL orders only against torch's current
        stream, so passing no stream (or the legacy default stream, handle 0) would let a collective start
        before the render that feeds it has finished: a real torch stream is required and made current."""
        import torch
        if stream is None:
            if getattr(self, "_stream", None) is None:
                self._stream = torch.cuda.Stream(device=self.backend.device)
            stream = self._stream
        if not hasattr(stream, "cuda_stream") or stream.cuda_stream == 0:
            raise ValueError("multi-GPU rendering needs a non-default torch.cuda.Stream (got the legacy default "
                             "stream or a raw handle): the render and NCCL must share one ordered stream")
        return stream

    def _partial(self, spp, depth, stream=None):
        """This rank's film as a torch tensor [h, w, 4] on its device (zeros where not rendered)."""
        import torch
        w, h = self.backend.size
        ptr = None
        cs = stream.cuda_stream if stream is not None else None
        if self.mode == "samples":
            for rows, samples, flags in self.plan(spp):
                ptr = self.backend.render_device(spp, depth, rows=rows, samples=samples, stream=cs, flags=flags)
        elif self.plan(spp):  # the interleaved bands of this rank in ONE call (pb2_render_params.row_band/parts/part)
            ptr = self.backend.render_device(spp, depth, stream=cs, bands=(self.band, self.world, self.rank))
        if ptr is None:  # nothing dealt to this rank (more ranks than bands/samples)
            return torch.zeros((h, w, 4), dtype=torch.float32, device=f"cuda:{self.backend.device}")

        class _P:
            __cuda_array_interface__ = {"shape": (h, w, 4), "typestr": "<f4", "data": (ptr, False),
                                        "version": 3, "strides": None}
        return torch.as_tensor(_P(), device=f"cuda:{self.backend.device}")

    def render(self, spp, depth, dst=0, stream=None):
        """-> the full film tensor on rank `dst` (other ranks get their partial back). Asynchronous on
        `stream` (a torch.cuda.Stream; one is created when None and world > 1): synchronise it before
        reading the film."""
        import contextlib
        st, ctx = None, contextlib.nullcontext()
        if self.backend.device != "cpu" and (self.world > 1 or stream is not None):
            import torch
            st = self._torch_stream(stream)
            ctx = torch.cuda.stream(st)
        with ctx:
            film = self._partial(spp, depth, st)
            if self.world > 1:
                import torch.distributed as dist
                dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=self.group)
        return film


def restage_sharded(backend, rank, world, stream=None, group=None):
    """Re-stage a scene that every rank of one NVLink domain holds a replica of: each rank pulls 1/world of
    every large array through PCIe from its pinned host copy (pb2_restage_scene_part) and the slices are
    all-gathered in place over NVLink / NVSwitch -- world x less host traffic per GPU than world full copies,
    which is what bounds the end-to-end step at 8 GPUs (1.43 GB per replica on C5). `stream`: the
    torch.cuda.Stream the copies AND the collectives are issued on (required when world > 1; the legacy
    default stream is rejected, see ShardedRenderer._torch_stream); the caller renders on the same stream."""
    if world == 1:
        backend.restage_part(rank, world, stream.cuda_stream if stream is not None else None)
        return 0
    import torch
    import torch.distributed as dist
    if stream is None or not hasattr(stream, "cuda_stream") or stream.cuda_stream == 0:
        raise ValueError("restage_sharded needs a non-default torch.cuda.Stream when world > 1")
    moved = 0
    with torch.cuda.stream(stream):
        arrays = backend.restage_part(rank, world, stream.cuda_stream)
        for ptr, nbytes, part in arrays:
            class _P:
                __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                            "version": 3, "strides": None}
            full = torch.as_tensor(_P(), device=f"cuda:{backend.device}")
            dist.all_gather_into_tensor(full, full[rank * part:(rank + 1) * part], group=group)
            moved += nbytes
    return moved
