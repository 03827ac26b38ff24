"""Worker for the multi-rank tests (launched by torch.distributed.run).

mode "gpu":  one rank per GPU; full biharmonic through libbwk_b200 with NCCL halo exchange;
             every rank saves its block of the result.
mode "cpu":  gloo only, no GPU: the ranks exchange exactly the rows the plan lists (the wire
             format of bwk's halo) and evaluate the plan's gather on the host.
"""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def hypervis_inputs(n, nlev):
    """dp, dpdiss (4,4,nlev,n) and dt*nu_q, the same on every rank and in the single-rank run of the test"""
    rng = np.random.default_rng(77)
    dp = np.asfortranarray(0.5 + rng.random((4, 4, nlev, n)))
    dpd = np.asfortranarray(0.5 + rng.random((4, 4, nlev, n)))
    return dp, dpd, 1.0e24  # with the LCG metric the weak biharmonic is O(1e-25): the two terms of the update are comparable


def user_elements(n):
    """element records (n, 144) that are NOT the generator's: the same on every rank and in the test"""
    return 0.25 + np.random.default_rng(79).random((n, 144))


def dp3d_v2c(n):
    """vec_sphere2cart stand-in (n, 96), the same on every rank and in the single-rank run of the test"""
    return np.random.default_rng(78).random((n, 96))


def main():
    mode, ne, nlev, qsize, outdir = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import codesign_kernels_b200 as bwk

    n = 6 * ne * ne
    first, count = bwk.partition(n, rank, world)
    if mode == "big":
        # BASELINE config 5 (ne512np4 L128 q10 on 8 GPUs, 32 GB of qtens per rank): device-generated data, full biharmonic
        # with the NCCL halo, twice (determinism); every rank saves a few of its elements for the patch oracle of the test
        import torch

        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        ku = bwk.KernelUtils(count, nlev, qsize).allocate_data(host_arrays=False)
        idb = [bwk.KernelUtils.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(idb, src=0)
        bwk.KernelUtils.comm_init(rank, world, idb[0])
        ku.initialize_data_device(n, first)
        ku.run_device(1, count)  # reference-pinned single Laplacian
        ku.sync()
        picks = [0, ne + 3, count // 2 + 7, count - 1]  # first row of the block (S neighbours on another rank), interior, last element
        np.save(os.path.join(outdir, f"big_lap_{rank}.npy"), np.stack([ku.download_elements(e, 1) for e in picks], axis=0))
        ku.set_topology_cubed_sphere(ne, first)
        sums = []
        for rep in range(2):
            ku.initialize_data_device(n, first)
            ku.run_biharmonic_device()
            ku.sync()
            sums.append(ku.checksum())
        assert sums[0] == sums[1], sums
        np.save(os.path.join(outdir, f"big_bih_{rank}.npy"), np.stack([ku.download_elements(e, 1) for e in picks], axis=0))
        np.save(os.path.join(outdir, f"big_picks_{rank}.npy"), np.array([first + e for e in picks], dtype=np.int64))
        dist.barrier()
        ku.finalize()
    elif mode == "dp3d":
        # biharmonic_wk_dp3d across ranks (context with qsize = 4: T, v, dp3d per level); device-generated numbers as fields
        import torch

        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        ku = bwk.KernelUtils(count, nlev, 4).allocate_data(host_arrays=False)
        idb = [bwk.KernelUtils.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(idb, src=0)
        bwk.KernelUtils.comm_init(rank, world, idb[0])
        ku.initialize_data_device(n, first)
        ku.set_topology_cubed_sphere(ne, first)
        ku.upload_vec_sphere2cart(np.ascontiguousarray(dp3d_v2c(n)[first : first + count]))
        for rep in range(2):  # twice: the halo buffers are reused across steps
            ku.initialize_data_device(n, first)
            ku.run_biharmonic_dp3d_device(rep == 1, True)
            ku.sync()
        np.save(os.path.join(outdir, f"dp3d_{rank}.npy"), ku.download_elements(0, count))
        dist.barrier()
        ku.finalize()
    elif mode == "gpu":
        import torch

        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        ku = bwk.KernelUtils(count, nlev, qsize).allocate_data(host_arrays=False)
        idb = [bwk.KernelUtils.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(idb, src=0)
        bwk.KernelUtils.comm_init(rank, world, idb[0])
        ku.initialize_data_device(n, first)
        ku.set_topology_cubed_sphere(ne, first)
        for rep in range(2):  # twice: the halo buffers are reused across steps
            ku.initialize_data_device(n, first)
            ku.run_biharmonic_device()
            ku.sync()
        np.save(os.path.join(outdir, f"dev_{rank}.npy"), ku.download_elements(0, count))
        np.save(os.path.join(outdir, f"bits_{rank}.npy"), np.array([ku.checksum_bits(first)], dtype=np.uint64))
        # advance_hypervis_scalar across ranks: two DSS exchanges per subcycle, halo buffers reused
        dp, dpd, coef = hypervis_inputs(n, nlev)
        ku.initialize_data_device(n, first)
        ku.upload_hypervis_fields(np.asfortranarray(dp[..., first : first + count]), np.asfortranarray(dpd[..., first : first + count]))
        ku.run_advance_hypervis_device(2.0, coef, 2, True)
        ku.sync()
        np.save(os.path.join(outdir, f"hv_{rank}.npy"), ku.download_elements(0, count))
        # caller-supplied element records on several ranks (a real mesh, not the generator): rspheremp = 1/DSS(spheremp) must
        # come out of the halo plan (the sharers' spheremp travels once with ncclSend/Recv), and the full biharmonic with it
        ku.elem = np.zeros(count, dtype=bwk.ELEMENT_DTYPE)
        ku.elem.view(np.float64).reshape(count, 144)[...] = user_elements(n)[first : first + count]
        ku.update_device(elem=True, qtens=False, deriv=False)
        np.save(os.path.join(outdir, f"rsph_user_{rank}.npy"), ku.get_rspheremp())
        ku.run_biharmonic_device()
        ku.sync()
        np.save(os.path.join(outdir, f"user_{rank}.npy"), ku.download_elements(0, count))
        dist.barrier()
        ku.finalize()
    else:
        g = bwk.cubed_sphere_gdof(ne)
        pl = bwk.Plan(n, g, rank, world)
        rng = np.random.default_rng(1234)
        field = rng.standard_normal(16 * n)  # same on every rank
        mine = np.full(16 * n, np.nan)
        mine[16 * first : 16 * (first + count)] = field[16 * first : 16 * (first + count)]

        def row_of(arr, e, row):
            s = arr[16 * e : 16 * e + 16]
            return s[4 * row : 4 * row + 4].copy() if row < 4 else s[(0 if row == 4 else 3) :: 4].copy()

        import torch

        reqs, recvbufs = [], {}
        for p in range(world):
            if p == rank:
                continue
            se, sr = pl.send_list(p)
            if se.size:
                buf = torch.from_numpy(np.concatenate([row_of(mine, int(e), int(r)) for e, r in zip(se, sr)]))
                reqs.append(dist.isend(buf, p))
            re_, rr = pl.recv_list(p)
            if re_.size:
                recvbufs[p] = torch.empty(4 * re_.size, dtype=torch.float64)
                reqs.append(dist.irecv(recvbufs[p], p))
        for r in reqs:
            r.wait()
        # scatter the received rows into the (otherwise NaN) global array: only those entries may be read
        for p, buf in recvbufs.items():
            re_, rr = pl.recv_list(p)
            b = buf.numpy().reshape(-1, 4)
            for k, (e, row) in enumerate(zip(re_, rr)):
                s = mine[16 * int(e) : 16 * int(e) + 16]
                if row < 4:
                    s[4 * row : 4 * row + 4] = b[k]
                else:
                    s[(0 if row == 4 else 3) :: 4] = b[k]
        out = pl.apply_host(mine)
        assert not np.isnan(out).any(), "the plan read a value that was neither local nor received"
        np.save(os.path.join(outdir, f"cpu_{rank}.npy"), out)
        if rank == 0:
            np.save(os.path.join(outdir, "field.npy"), field)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
