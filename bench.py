#!/usr/bin/env python
"""bench.py — headline benchmark: Jacobi3D cell updates per second, 512^3 fp32 cells per GPU, weak scaling, plus the
other two parts of BASELINE.json's metric ("Exchange us/iter + Jacobi3D/Astaroth Gcell/s") on their named configs.

  python bench.py --gpus 1 --steps 30 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P bench.py --gpus 8 ...
  python bench.py --impl reference ...      (the unmodified reference's own binaries, see baseline/)

One JSON line on rank 0. What the own arm does, in order:

  verify    at the bench size on the bench's GPUs: a random initial field (a function of the global coordinate), 20
            iterations with the plain kernel + exchange() schedule (Mode NoOverlap), the same with the fused compute +
            NVLink-push kernel, and the two 64-bit field fingerprints must be equal ("verified": true means every cell
            of every GPU agreed bit for bit). Not timed.
  value     K fused iterations (compute of the whole region + halo push + flags; nothing skipped — this workload has no
            optimizer-like extra step), bracketed by barrier + device synchronisation, CUDA events per GPU, max over
            GPUs and ranks; repeated R times, the MEDIAN repeat is reported. The field (512 MB) and its double buffer
            are far larger than the 126 MB L2, so every iteration streams from HBM.
  e2e       the same through the public Python API with, every step, a pinned-host -> device upload of the step's inputs
            (sphere temperatures) and a device -> pinned-host download of the step's result (temperature at each
            subdomain centre), host wall clock between barriers, max over ranks; median of R repeats.
  exchange_us_per_iter   the `weak` config (radius 3, 4 x fp32, 26 directions, 512^3 cells per GPU): K blocking
            DistributedDomain.exchange() calls, device-timed (events on every exchange stream, max over GPUs and ranks).
  astaroth_gcell_per_s   BASELINE config 3 (8 x fp64 fields, radius 3, all 26 halo directions, 256^3 cells per GPU, weak):
            K AstarothSim steps, device-timed the same way."""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "jacobi3d_gcell_updates_per_s"


def weak_size(per_gpu, n):
    """The reference's weak scaling rule (bin/jacobi3d.cu:167-169): round(per_gpu * n^0.33333) per axis."""
    return int(float(per_gpu) * pow(float(n), 0.33333) + 0.5)


def shared_config(size, n):
    """The part of the JSON line both arms must print identically."""
    X = weak_size(size, n)
    return {
        "model": "jacobi3d 7-point radius-1",
        "global_batch": f"{X}x{X}x{X} cells",
        "seq_len": f"{size}^3 cells per GPU",
        "parallelism": f"dd{n}",
        "l2": "inputs larger than L2 (512 MB field per GPU vs 126 MB L2); no flush needed",
    }


def run_own(args):
    import ctypes as C
    import numpy as np
    from stencil2_b200._lib import lib
    from stencil2_b200.parallel import comm
    from stencil2_b200.models import Jacobi3D, AstarothSim, DistributedDomain, Method
    from stencil2_b200.utils.clocks import ClockSampler

    os.environ.setdefault("STENCIL_LOG_LEVEL", "2")
    rank, world = comm.init_from_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n = args.gpus
    if world == 1 and n > 1:
        gpus = list(range(n))  # one process driving n GPUs (the reference's default deployment)
        deployment = f"1 process x {n} GPUs"
    else:
        gpus = [local_rank] if world > 1 else [0]
        deployment = f"{world} processes x 1 GPU" if world > 1 else "1 process x 1 GPU"
    X = weak_size(args.size, n)
    L = lib()
    K, W, R = args.steps, max(args.warmup, 3), max(args.repeats, 1)
    cells = float(X) ** 3
    nproc = world if world > 1 else 1

    # ---- verification: fused == plain kernel + exchange(), bit for bit, at the bench size on these GPUs ----
    verified, fp_ref, fp_fused = None, None, None
    if args.verify_iters > 0:
        jn = Jacobi3D(X, X, X, methods=Method.ALL, mode=Jacobi3D.NO_OVERLAP, gpus=gpus, tile_z=args.tile_z)
        jn.init_random(1)
        for _ in range(args.verify_iters):
            jn.step()
        fp_ref = jn.fingerprint()
        jn.close()
        comm.barrier()
    j = Jacobi3D(X, X, X, methods=Method.ALL, mode=Jacobi3D.FUSED, gpus=gpus, tile_z=args.tile_z)
    mode = {0: "fused", 1: "overlap", 2: "no-overlap"}[j.mode]
    grid = j.domain.grid
    if args.verify_iters > 0:
        j.init_random(1)
        for _ in range(args.verify_iters):
            j.step()
        fp_fused = j.fingerprint()
        verified = bool(fp_fused == fp_ref)
    if args.random_init:
        j.init_random(2)
    else:
        j.init()

    # ---- device-timed kernel path ----
    j.run(W)
    # clocks and throttle reasons are sampled every millisecond from here to the end of the end-to-end region
    sampler = ClockSampler(local_rank, period_s=0.001).start() if rank == 0 else None
    l0 = j.launches
    secs_all = [j.run(K) for _ in range(R)]
    launches = (j.launches - l0) // R
    secs = statistics.median(secs_all)
    value = cells * K / secs / 1e9

    # ---- end-to-end through the public API: per-step pinned H2D inputs + D2H result ----
    nsub = len(gpus)
    params = L.stencil_host_alloc_pinned(64)
    results = L.stencil_host_alloc_pinned(4 * nsub * K)
    pf = (C.c_float * 16).from_address(params)
    pf[0], pf[1] = 1.0, 0.0
    for _ in range(3):
        j.step_io(params, 8, results)
    j.sync()
    e2e_all, finite = [], True
    for _ in range(R):
        comm.barrier()
        t0 = time.perf_counter()
        for s in range(K):
            pf[0] = 1.0  # this step's inputs are written to pinned memory by the host, then uploaded
            j.step_io(params, 8, results + 4 * nsub * s)
        j.sync()
        e2e_all.append(comm.allreduce_max(time.perf_counter() - t0))
        res = (C.c_float * (nsub * K)).from_address(results)
        finite = finite and all(0.0 <= res[i] <= 1.0 for i in range(nsub * K))
    clocks = sampler.stop() if sampler else None
    e2e_secs = statistics.median(e2e_all)
    e2e_value = cells * K / e2e_secs / 1e9
    checksum = j.checksum()
    L.stencil_host_free_pinned(params)
    L.stencil_host_free_pinned(results)
    j.close()
    comm.barrier()

    # ---- halo exchange alone: the `weak` config (radius 3, 4 x fp32, 512^3 per GPU) ----
    exchange = None
    if not args.skip_exchange:
        try:
            dd = DistributedDomain(X, X, X)
            dd.set_gpus(gpus)
            dd.set_radius(3)
            dd.set_methods(Method.ALL)
            dd.set_plan_file(False)
            for q in range(4):
                dd.add_data(np.float32, f"d{q}")
            dd.realize()
            dd.time_exchange(3)
            host_s, dev_s = dd.time_exchange(K)
            exchange = {"us_per_exchange_device": round(dev_s / K * 1e6, 3), "us_per_exchange_host": round(host_s / K * 1e6, 3),
                        "bytes_per_exchange": int(dd.exchange_bytes_for_method(Method.ALL)) * 1,
                        "config": f"radius 3, 4 x fp32, 26 directions, {X}^3 cells, {args.size}^3 per GPU", "grid": list(dd.grid)}
            dd.close()
        except Exception as e:  # noqa: BLE001 - the headline above must still be printed
            exchange = {"us_per_exchange_device": None, "error": repr(e)[:300]}
        comm.barrier()

    # ---- Astaroth-style workload: BASELINE config 3 ----
    astaroth = None
    if not args.skip_astaroth:
        try:
            A = weak_size(args.astaroth_size, n)
            sim = AstarothSim(A, A, A, fields=8, fp64=True, stencil=AstarothSim.STAR19, gpus=gpus)
            sim.run(W)
            a_all = [sim.run_timed(K) for _ in range(R)]
            a_dev = statistics.median([t[2] for t in a_all])
            a_host = statistics.median([t[0] for t in a_all])
            astaroth = {"gcell_per_s": round(float(A) ** 3 * K / a_dev / 1e9, 3), "ms_per_step_device": round(a_dev / K * 1e3, 5),
                        "ms_per_step_host": round(a_host / K * 1e3, 5), "launches_per_step": int(sim.launches // (W + R * K + 1)),
                        "config": f"8 x fp64 fields, radius 3, 19-point 6th-order step, all 26 halo directions kept current, {A}^3 cells, "
                                  f"{args.astaroth_size}^3 per GPU", "grid": list(sim.domain.grid)}
            sim.close()
        except Exception as e:  # noqa: BLE001
            astaroth = {"gcell_per_s": None, "error": repr(e)[:300]}
        comm.barrier()

    if rank == 0:
        out = {
            "metric": METRIC,
            "value": round(value, 3),
            "unit": "Gcell/s",
            "n_gpus": n,
            "steps": K,
            "warmup": W,
            "repeats": R,
            "ms_per_step": round(secs / K * 1e3, 5),
            "ms_per_step_repeats": [round(s / K * 1e3, 5) for s in secs_all],
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "fp32",
            "data": ("synthetic (random initial field, a function of the global coordinate, plus hot/cold Dirichlet spheres)" if args.random_init
                     else "synthetic (field initialised to 0.5 with hot/cold Dirichlet spheres, as the reference app)"),
            "impl": "own",
            "config": shared_config(args.size, n),
            "subdomain_grid": list(grid),
            "deployment": f"{deployment}, {mode} compute+halo-push kernel",
            "timing": "CUDA events per GPU around the K steps, max over GPUs and ranks, median of R repeats",
            "verified": verified,
            "verify": {"iters": args.verify_iters, "fingerprint_fused": f"{fp_fused:016x}" if fp_fused is not None else None,
                       "fingerprint_no_overlap": f"{fp_ref:016x}" if fp_ref is not None else None,
                       "what": "random initial field; fused kernel vs plain kernel + exchange(); equal fingerprints = every cell equal bit for bit"},
            "clocks": clocks,
            "e2e": {
                "value": round(e2e_value, 3),
                "unit": "Gcell/s",
                "ms_per_step": round(e2e_secs / K * 1e3, 5),
                "h2d_bytes_per_step": 8 * nsub * nproc,
                "d2h_bytes_per_step": 4 * nsub * nproc,
                "results_valid": bool(finite),
                "timing": "host wall clock between barrier+sync points, max over ranks, median of R repeats",
            },
            "gpu_launches": int(launches * nproc),
            "exchange_us_per_iter": exchange["us_per_exchange_device"] if exchange else None,
            "exchange": exchange,
            "astaroth_gcell_per_s": astaroth["gcell_per_s"] if astaroth else None,
            "astaroth": astaroth,
            "checksum": checksum,
            "hbm_fraction_of_measured": None,
        }
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peak = json.load(f)["hbm_gbs"]
            out["hbm_fraction_of_measured"] = round((8.0 * float(args.size) ** 3 / (secs / K)) / 1e9 / peak, 4)
        except Exception:
            pass
        print(json.dumps(out), flush=True)
    comm.finalize()


def run_reference(args):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import reference_arm
    reference_arm.main(args, shared_config(args.size, args.gpus), METRIC)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--repeats", type=int, default=5, help="the K timed steps are repeated this many times; the median is reported")
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--size", type=int, default=512, help="cells per axis per GPU")
    ap.add_argument("--astaroth-size", type=int, default=256, help="cells per axis per GPU of the Astaroth config")
    ap.add_argument("--tile-z", type=int, default=0, help="planes per tile; 0 = the library's own choice")
    ap.add_argument("--verify-iters", type=int, default=20, help="0 skips the fused-vs-plain verification")
    ap.add_argument("--random-init", action="store_true", help="time on a random initial field instead of the reference app's 0.5")
    ap.add_argument("--skip-exchange", action="store_true")
    ap.add_argument("--skip-astaroth", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
