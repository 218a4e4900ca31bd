"""Every GPU of the box behind ONE render call, inside the library (include/pine_b200.h pb2_multi_*): the
north_star's "scene replicated, pixel tiles or sample passes sharded, a single NCCL reduce of the float film".
The single-device cases (a pb2_multi over one GPU, the interleaved row bands of pb2_render_params) run on any
box; the N > 1 cases are skipped below two GPUs (gpurun --gpus 2). Reference interface the drop-in stays behind:
Integrator::render(Scene&), src/pine/core/integrator.h:20-25."""
import numpy as np
import pytest

import common
import pine_b200 as pb
from common import relmse, u32
from conftest import needs_ref
from pine_b200 import _abi as A
from pine_b200 import scenes

pytestmark = pytest.mark.gpu


def _n_gpus():
    return A.lib().pb2_device_count()


two_gpus = pytest.mark.skipif(_n_gpus() < 2, reason="needs two CUDA devices")


def _scene(size=(96, 80)):
    S = pb.Scene()
    cfg = scenes.cbox_readme(S, size=size)
    return S, cfg


def test_multi_over_one_device_is_the_single_device_render():
    S, cfg = _scene()
    B = pb.Backend(0)
    B.upload_scene(S)
    want = B.render(8, cfg["depth"])
    M = pb.MultiBackend(devices=[0])
    assert M.n == 1
    M.upload_scene(S)
    for mode in ("rows", "samples"):
        assert np.array_equal(u32(M.render(8, cfg["depth"], mode=mode)), u32(want))
    M.restage()
    seen = []
    M.set_progress_callback(seen.append)   # pb2_multi_set_progress_callback: mean of the devices' fractions
    assert np.array_equal(u32(M.render(8, cfg["depth"])), u32(want))
    assert seen and seen[-1] == 1.0 and seen == sorted(seen)
    M.set_progress_callback(None)
    c = M.counters()
    assert c["last_closest_rays"] == B.counters()["last_closest_rays"]
    M.close()
    B.close()


@pytest.mark.parametrize("height,band,parts", [(80, 16, 3), (77, 16, 4), (50, 8, 2), (10, 16, 8)])
def test_interleaved_row_bands_assemble_the_full_film(height, band, parts):
    """pb2_render_params.row_band / row_parts / row_part on ONE device: the parts render disjoint rows (the
    rest stays zero), and accumulated with PB2_RENDER_KEEP_FILM they give the full film bit for bit."""
    S, cfg = _scene(size=(64, height))
    B = pb.Backend(0)
    B.upload_scene(S)
    want = B.render(4, cfg["depth"])
    owner = (np.arange(height) // band) % parts
    for part in range(parts):
        f = B.render(4, cfg["depth"], bands=(band, parts, part))
        assert np.array_equal(u32(f[owner == part]), u32(want[owner == part]))
        assert (f[owner != part] == 0).all()
    for part in range(parts):
        acc = B.render(4, cfg["depth"], bands=(band, parts, part), flags=4 if part else 0)
    assert np.array_equal(u32(acc), u32(want))
    with pytest.raises(pb.PineB200Error):
        B.render(4, cfg["depth"], bands=(0, 2, 0))
    with pytest.raises(pb.PineB200Error):
        B.render(4, cfg["depth"], bands=(16, 2, 2))
    B.close()


@two_gpus
def test_two_gpus_rows_are_bit_identical_and_samples_statistically_equal():
    S, cfg = _scene(size=(160, 120))
    B = pb.Backend(0)
    B.upload_scene(S)
    spp, depth = 16, cfg["depth"]
    want = B.render(spp, depth)
    M = pb.MultiBackend()
    assert M.n == _n_gpus() >= 2
    M.upload_scene(S)
    rows = M.render(spp, depth, mode="rows")
    assert np.array_equal(u32(rows), u32(want))
    c = M.counters()
    assert c["last_closest_rays"] == B.counters()["last_closest_rays"]
    samples = M.render(spp, depth, mode="samples")
    assert np.allclose(samples[..., 3], 1.0)
    # each GPU's share is the single-GPU render of that sample range (reseeded streams): sum them on one GPU
    from pine_b200.multigpu import sample_shards
    acc = sum(B.render(spp, depth, samples=s).astype(np.float64) for s in sample_shards(spp, M.n))
    np.testing.assert_allclose(samples, acc, rtol=1e-5, atol=1e-6)
    # a changed host scene reaches EVERY replica through the sharded staging + all-gather
    S2 = pb.Scene()
    scenes.cbox_indirect(S2, size=(160, 120))
    M.upload_scene(S2)
    B.upload_scene(S2)
    assert np.array_equal(u32(M.render(spp, depth, mode="rows")), u32(B.render(spp, depth)))
    M.close()
    B.close()


@two_gpus
def test_two_gpus_report_progress_of_the_whole_job():
    if A.lib().pb2_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    S, cfg = _scene()
    M = pb.MultiBackend(devices=[0, 1])
    M.upload_scene(S)
    seen = []
    M.set_progress_callback(seen.append)
    for mode in ("rows", "samples"):
        del seen[:]
        M.render(8, cfg["depth"], mode=mode)
        assert len(seen) >= 2 and max(seen) == 1.0 and min(seen) > 0.0, (mode, seen)
    M.close()


@two_gpus
@pytest.mark.parametrize("fn", ["classic", "plane_yard"])
def test_two_gpus_carry_node_programs_planes_and_cylinders(fn):
    """The round-2 scene content reaches every replica: node programs (and the switch to the generic shade kernels
    they need), Plane / Cylinder records (and the traversal kernels compiled with them): rows over all GPUs equal
    the single-GPU film bit for bit."""
    S = pb.Scene()
    cfg = scenes.SCENES[fn](S, size=(96, 64))
    B = pb.Backend(0)
    B.upload_scene(S)
    want = B.render(8, cfg["depth"])
    M = pb.MultiBackend()
    M.upload_scene(S)
    assert np.array_equal(u32(M.render(8, cfg["depth"], mode="rows")), u32(want))
    M.restage()
    assert np.array_equal(u32(M.render(8, cfg["depth"], mode="rows")), u32(want))
    M.close()
    B.close()


@two_gpus
def test_two_gpus_large_mesh_replicas_are_complete():
    """A scene whose arrays exceed the sharded-staging threshold (1 MiB): 1/N through PCIe + in-place all-gather
    must leave the same BVH on every GPU -- rows mode is then bit-identical to one GPU."""
    S = pb.Scene()
    cfg = scenes.mesh_field(S, n_instances=64, subdiv=4, seed=7, size=(256, 144), spp=8, depth=5, with_shapes=True)
    B = pb.Backend(0)
    B.upload_scene(S)
    want = B.render(8, 5)
    M = pb.MultiBackend()
    M.upload_scene(S)
    assert np.array_equal(u32(M.render(8, 5, mode="rows")), u32(want))
    M.restage()
    assert np.array_equal(u32(M.render(8, 5, mode="rows")), u32(want))
    M.close()
    B.close()


@two_gpus
@needs_ref
def test_dropin_shim_uses_every_gpu_and_matches_one_gpu():
    """One live pine::Scene through Integrator::render(Scene&) -> pine::CudaPathIntegrator with its default
    device list (every visible GPU, row bands): the film equals the 1-GPU film bit for bit."""
    from oracle import ref
    R = ref.RefScene()
    cfg = scenes.cbox_readme(R, size=(160, 160))
    one, _, st1 = R.render_cuda(16, cfg["depth"], device=0)
    allg, _, stn = R.render_cuda(16, cfg["depth"], device=-1)
    assert st1["n_devices"] == 1 and stn["n_devices"] == _n_gpus()
    assert np.array_equal(u32(one), u32(allg))
    assert stn["closest_rays"] == st1["closest_rays"]
    film_ref, _ = R.render(16, cfg["depth"])
    assert relmse(allg, film_ref) <= 1e-4


def test_imported_scene_renders_like_the_built_one(tmp_path):
    """One host build per node for one-process-per-GPU setups: a context that IMPORTS the scene another context
    built (pb2_export_host_scene / pb2_import_host_scene) renders the same film bits, before and after a
    re-staging from the mapped file, and its sharded staging slices are usable (bench.py at N > 1)."""
    S = pb.Scene()
    cfg = scenes.mesh_field(S, n_instances=64, subdiv=3, seed=11, size=(160, 96), spp=8, depth=5, with_shapes=True)
    A = pb.Backend(0)
    A.upload_scene(S)
    want = A.render(8, 5)
    path = str(tmp_path / "scene.bin")
    A.export_scene(path)
    B = pb.Backend(0)
    B.import_scene(path, S.size)
    import os
    os.unlink(path)  # the mapping keeps the pages
    assert np.array_equal(u32(B.render(8, 5)), u32(want))
    B.restage()
    assert np.array_equal(u32(B.render(8, 5)), u32(want))
    for part in range(2):  # both halves of the sharded staging, one after the other = the whole scene again
        B.restage_part(part, 2)
    assert np.array_equal(u32(B.render(8, 5)), u32(want))
    rays = A.primary_rays(0)
    ha, hb = A.trace_closest(rays), B.trace_closest(rays)
    assert np.array_equal(ha["prim"], hb["prim"]) and np.array_equal(u32(ha["t"]), u32(hb["t"]))
    A.close()
    B.close()


def _torch_rank(rank, world, port, out_dir):
    """One rank of the one-process-per-GPU path (pine_b200/multigpu.py) over NCCL."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import torch.distributed as dist
    import pine_b200 as pb2
    from pine_b200 import scenes as sc
    from pine_b200.multigpu import ShardedRenderer, restage_sharded
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    B = pb2.Backend(rank)
    ts = torch.cuda.Stream(device=rank)
    for tag, fn in (("a", sc.cbox_readme), ("b", sc.cbox_indirect)):  # the host scene CHANGES between the stages
        S = pb2.Scene()
        fn(S, size=(160, 128))
        B.upload_scene(S)
        R = ShardedRenderer(B, rank, world, mode="rows", band=16)
        for rep in range(2):
            restage_sharded(B, rank, world, ts)  # 1/world through PCIe + in-place all-gather, same stream as the render
            film = R.render(8, 4, dst=0, stream=ts)
            ts.synchronize()
            if rank == 0:
                np.save(os.path.join(out_dir, f"film_{tag}{rep}.npy"), film.cpu().numpy())
    try:  # the legacy default stream must be refused (NCCL would not be ordered after the render)
        restage_sharded(B, rank, world, None)
        raise SystemExit("restage_sharded accepted no stream")
    except ValueError:
        pass
    dist.barrier()
    dist.destroy_process_group()


@two_gpus
def test_torch_ranks_over_nccl_restage_and_render_in_stream_order(tmp_path):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_torch_rank, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    B = pb.Backend(0)
    for tag, fn in (("a", scenes.cbox_readme), ("b", scenes.cbox_indirect)):
        S = pb.Scene()
        fn(S, size=(160, 128))
        B.upload_scene(S)
        want = B.render(8, 4)
        for rep in range(2):
            got = np.load(str(tmp_path / f"film_{tag}{rep}.npy"))
            assert np.array_equal(u32(got), u32(want)), (tag, rep)
    B.close()
