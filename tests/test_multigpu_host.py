"""Host-side logic of the multi-GPU path (pine_b200/multigpu.py) on CPU: the shard plans, and the
world_size-2 flow over the gloo backend with a CPU film model standing in for the per-GPU Backend
(the kernels are covered by the -m gpu suite; the N>1 GPU flow by bench.py under torchrun)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sample_shards_cover_and_balance():
    from pine_b200.multigpu import sample_shards
    for spp in (1, 2, 16, 64, 256):
        for world in (1, 2, 3, 4, 8):
            s = sample_shards(spp, world)
            assert len(s) == world and s[0][0] == 0 and s[-1][1] == spp
            assert all(a[1] == b[0] for a, b in zip(s, s[1:]))
            sizes = [e - b for b, e in s]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sample_shards(0, 2)


def test_row_bands_partition_the_film():
    from pine_b200.multigpu import row_bands
    for h in (1, 15, 16, 17, 640, 1080, 2160):
        for world in (1, 2, 4, 8):
            seen = np.zeros(h, np.int32)
            for r in range(world):
                for b, e in row_bands(h, world, r, band=16):
                    assert 0 <= b < e <= h and e - b <= 16
                    seen[b:e] += 1
            assert (seen == 1).all()
    # ragged: more ranks than bands leaves some ranks empty, never overlapping
    assert row_bands(10, 8, 3, band=16) == []


def _model_film(w, h, spp, rows, samples):
    """A deterministic stand-in for Backend.render: film[y, x] = sum_s g(x, y, s) / spp over the
    samples of the range, alpha = n/spp, zero outside the rows -- the contract of pb2_render_device
    (include/pine_b200.h: pb2_render_params)."""
    import torch
    y = torch.arange(h, dtype=torch.float32).view(h, 1, 1)
    x = torch.arange(w, dtype=torch.float32).view(1, w, 1)
    s = torch.arange(samples[0], samples[1], dtype=torch.float32).view(1, 1, -1)
    g = torch.sin(x * 0.37 + y * 0.11 + s * 1.3) ** 2 + 0.25
    film = torch.zeros(h, w, 4)
    val = g.sum(-1) / spp
    film[rows[0]:rows[1], :, 0] = val[rows[0]:rows[1]]
    film[rows[0]:rows[1], :, 1] = 2 * val[rows[0]:rows[1]]
    film[rows[0]:rows[1], :, 2] = 0.5 * val[rows[0]:rows[1]]
    film[rows[0]:rows[1], :, 3] = (samples[1] - samples[0]) / spp
    return film


def _worker(rank, world, port, mode, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from pine_b200.multigpu import PB2_RENDER_KEEP_FILM, ShardedRenderer
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)

    class ModelBackend:
        size, device = (40, 50), "cpu"

    class Model(ShardedRenderer):
        def _partial(self, spp, depth, stream=None):
            w, h = self.backend.size
            film = torch.zeros(h, w, 4)
            for i, (rows, samples, flags) in enumerate(self.plan(spp)):
                assert bool(flags & PB2_RENDER_KEEP_FILM) == (i > 0)
                part = _model_film(w, h, spp, rows, samples)
                film = film + part if flags & PB2_RENDER_KEEP_FILM else part
            return film

    R = Model(ModelBackend(), rank, world, mode=mode, band=16)
    film = R.render(16, 4, dst=0)
    if rank == 0:
        np.save(out_path, film.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["samples", "rows"])
def test_two_ranks_over_gloo_assemble_the_full_film(tmp_path, mode):
    import torch.multiprocessing as mp
    out = str(tmp_path / f"film_{mode}.npy")
    for attempt in range(3):  # the free port found here can be taken before the ranks bind it
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        try:
            mp.spawn(_worker, args=(2, port, mode, out), nprocs=2, join=True)
            break
        except Exception:
            if attempt == 2:
                raise
    got = np.load(out)
    full = _model_film(40, 50, 16, (0, 50), (0, 16)).numpy()
    assert (got[..., 3] == 1).all()
    if mode == "rows":
        assert np.array_equal(got, full)           # disjoint writers: bit-identical to one rank
    else:
        np.testing.assert_allclose(got, full, rtol=1e-6, atol=1e-7)  # fp32 sum order differs
