"""Multi-rank paths.  CPU (gloo, world_size 2 and 3): the halo wire lists carry exactly what the
gather needs.  GPU (-m gpu, needs >= 2 GPUs): N-rank result is bit-identical to the 1-rank result
and matches the oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest

from patch_oracle import patch_oracle_full as _patch_oracle_full

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mg_worker.py")


def _launch(nproc, args, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world,ne", [(2, 3), (3, 4)])
def test_gloo_halo_exchange_feeds_the_gather(bwk, oracle, tmp_path, world, ne):
    _launch(world, ["cpu", ne, 1, 1, tmp_path], 29511 + world)
    n = 6 * ne * ne
    field = np.load(tmp_path / "field.npy")
    ref = oracle.Dss(n, oracle.cubed_sphere_gdof(ne)).apply(np.asfortranarray(field.reshape((4, 4, 1, 1, n), order="F")).copy(order="F"))
    got = np.concatenate([np.load(tmp_path / f"cpu_{r}.npy") for r in range(world)])
    assert np.array_equal(got, ref.reshape(-1, order="F"))


@pytest.mark.gpu
@pytest.mark.parametrize("ne,nlev,qsize", [(4, 8, 2), (8, 72, 3)])
def test_multi_gpu_matches_single_gpu_bitwise(bwk, oracle, tmp_path, ne, nlev, qsize):
    import torch

    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if ngpu < 4 else 4
    n = 6 * ne * ne
    _launch(world, ["gpu", ne, nlev, qsize, tmp_path], 29533)
    multi = np.concatenate([np.load(tmp_path / f"dev_{r}.npy") for r in range(world)], axis=-1)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.run_biharmonic_device()
        single = ku.download_elements(0, n)
        single_bits = ku.checksum_bits(0)
    assert np.array_equal(multi, single)
    # the partition-independent bit checksum bench.py prints for every N: the per-rank values add up (mod 2^64) to the 1-rank value
    with np.errstate(over="ignore"):
        assert int(np.sum([np.load(tmp_path / f"bits_{r}.npy")[0] for r in range(world)], dtype=np.uint64)) == single_bits
    # advance_hypervis_scalar: N ranks == 1 rank bit for bit
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from mg_worker import hypervis_inputs

    dp, dpd, coef = hypervis_inputs(n, nlev)
    multi_hv = np.concatenate([np.load(tmp_path / f"hv_{r}.npy") for r in range(world)], axis=-1)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.upload_hypervis_fields(dp, dpd)
        ku.run_advance_hypervis_device(2.0, coef, 2, True)
        single_hv = ku.download_elements(0, n)
    assert np.array_equal(multi_hv, single_hv)
    dvv, elem, q = oracle.initialize_data(n, nlev, qsize)
    dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
    ref = dss.biharmonic_full(elem, dss.rspheremp(elem), q, dvv)
    assert oracle.max_rel(np.asfortranarray(multi), ref) <= 1e-12
    # caller-supplied element records: rspheremp through the halo plan == the oracle's 1/DSS(spheremp) bit for bit,
    # and the full biharmonic with those elements matches the oracle (q after the hypervis call above is what the device held:
    # the worker did not re-initialise it, so compare through the 1-rank device run instead)
    from mg_worker import user_elements

    uel = user_elements(n)
    rs_multi = np.concatenate([np.load(tmp_path / f"rsph_user_{r}.npy") for r in range(world)], axis=-1)
    assert np.array_equal(rs_multi, dss.rspheremp(uel))
    multi_user = np.concatenate([np.load(tmp_path / f"user_{r}.npy") for r in range(world)], axis=-1)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.upload_hypervis_fields(dp, dpd)
        ku.run_advance_hypervis_device(2.0, coef, 2, True)  # the field the ranks held when the user elements arrived
        ku.elem = np.zeros(n, dtype=bwk.ELEMENT_DTYPE)
        ku.elem.view(np.float64).reshape(n, 144)[...] = uel
        ku.update_device(elem=True, qtens=False, deriv=False)
        ku.run_biharmonic_device()
        single_user = ku.download_elements(0, n)
    assert np.array_equal(multi_user, single_user)


@pytest.mark.gpu
def test_dp3d_multi_gpu_matches_single_gpu_bitwise(bwk, tmp_path):
    """biharmonic_wk_dp3d: N ranks == 1 rank bit for bit (the element-local sweeps run boundary elements first and the DSS
    strip sweep reads the NCCL halo, as the scalar path does)"""
    import torch

    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if ngpu < 4 else 4
    ne, nlev = 6, 40
    n = 6 * ne * ne
    _launch(world, ["dp3d", ne, nlev, 4, tmp_path], 29537)
    multi = np.concatenate([np.load(tmp_path / f"dp3d_{r}.npy") for r in range(world)], axis=-1)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from mg_worker import dp3d_v2c

    with bwk.KernelUtils(n, nlev, 4).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.upload_vec_sphere2cart(dp3d_v2c(n))
        ku.run_biharmonic_dp3d_device(True, True)
        single = ku.download_elements(0, n)
    assert np.array_equal(multi, single)


@pytest.mark.gpu
def test_ne512_l128_q10_on_8_gpus(bwk, oracle, tmp_path):
    """BASELINE config 5: ne512np4 nlev=128 qsize=10 (1,572,864 elements, 257.7 GB of qtens) on 8 GPUs.  The single
    Laplacian and the full biharmonic of elements at block edges (halo path) and in the interior are compared with
    the oracle evaluated on the element's own neighbourhood; two runs give bit-identical checksums."""
    import torch

    if torch.cuda.device_count() < 8:
        pytest.skip("needs 8 GPUs")
    ne, nlev, qsize, world = 512, 128, 10, 8
    n = 6 * ne * ne
    _launch(world, ["big", ne, nlev, qsize, tmp_path], 29541)
    gd = oracle.cubed_sphere_gdof(ne).reshape(n, 16)
    for r in (0, 3, 7):
        picks = np.load(tmp_path / f"big_picks_{r}.npy")
        lap = np.load(tmp_path / f"big_lap_{r}.npy")
        bih = np.load(tmp_path / f"big_bih_{r}.npy")
        for k, e in enumerate(picks):
            dvv, elem, q = oracle.initialize_subset(n, nlev, qsize, int(e), 1)
            ref = oracle.biharmonic_wk_scalar_cpu(elem, q, dvv)
            assert oracle.max_rel(np.asfortranarray(lap[k]), ref) <= 1e-12
            if k in (0, 2):
                ref2 = _patch_oracle_full(oracle, ne, nlev, qsize, gd, int(e))
                assert oracle.max_rel(np.asfortranarray(bih[k]), ref2) <= 1e-12, (r, int(e))
