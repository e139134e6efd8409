#!/usr/bin/env python
"""bench.py — headline benchmark: Jacobi3D cell updates per second, 512^3 fp32 cells per GPU, weak scaling.

BASELINE.json metric "Exchange us/iter + Jacobi3D/Astaroth Gcell/s, weak/strong @1/2/4/8 B200", config
"jacobi3d 7-point radius-1 fp32, 512^3 cells per GPU weak scaling". One JSON line on rank 0.

  python bench.py --gpus 1 --steps 30 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P bench.py --gpus 8 ...
  python bench.py --impl reference ...      (the unmodified reference's own jacobi3d binary, see baseline/)

What is timed (own arm): K fused iterations (compute of the whole compute region + halo push + flags; nothing is
skipped, there is no optimizer-like extra step in this workload) bracketed by barrier + device synchronisation,
CUDA events per GPU, max over ranks. The field (512 MB) and its double buffer are far larger than the 126 MB L2,
so every iteration streams from HBM. `e2e` repeats the measurement through the public Python API with, every step, a
pinned-host -> device upload of the step's inputs (sphere temperatures) and a device -> pinned-host download of the
step's result (temperature at each subdomain centre), wall-clock timed between barriers, max over ranks."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def weak_size(per_gpu, n):
    """The reference's weak scaling rule (bin/jacobi3d.cu:167-169): round(per_gpu * n^0.33333) per axis."""
    return int(float(per_gpu) * pow(float(n), 0.33333) + 0.5)


def run_own(args):
    import ctypes as C
    from stencil2_b200._lib import lib
    from stencil2_b200.parallel import comm
    from stencil2_b200.models import Jacobi3D, Method
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
    j = Jacobi3D(X, X, X, methods=Method.ALL, mode=Jacobi3D.FUSED, gpus=gpus, tile_z=args.tile_z)
    mode = {0: "fused", 1: "overlap", 2: "no-overlap"}[j.mode]
    cells = float(X) ** 3

    # ---- device-timed kernel path ----
    j.run(max(args.warmup, 3))
    # clocks and throttle reasons are sampled every millisecond from here to the end of the end-to-end region
    sampler = ClockSampler(local_rank, period_s=0.001).start() if rank == 0 else None
    l0 = j.launches
    secs = j.run(args.steps)
    launches = j.launches - l0
    value = cells * args.steps / secs / 1e9

    # ---- end-to-end through the public API: per-step pinned H2D inputs + D2H result ----
    nsub = len(gpus)
    params = L.stencil_host_alloc_pinned(64)
    results = L.stencil_host_alloc_pinned(4 * nsub * args.steps)
    pf = (C.c_float * 16).from_address(params)
    pf[0], pf[1] = 1.0, 0.0
    for _ in range(3):
        j.step_io(params, 8, results)
    j.sync()
    comm.barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        pf[0] = 1.0  # this step's inputs are written to pinned memory by the host, then uploaded
        j.step_io(params, 8, results + 4 * nsub * s)
    j.sync()
    e2e_secs = comm.allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop() if sampler else None
    res = (C.c_float * (nsub * args.steps)).from_address(results)
    finite = all(0.0 <= res[i] <= 1.0 for i in range(nsub * args.steps))
    e2e_value = cells * args.steps / e2e_secs / 1e9
    checksum = j.checksum()
    L.stencil_host_free_pinned(params)
    L.stencil_host_free_pinned(results)
    j.close()
    comm.barrier()

    if rank == 0:
        out = {
            "metric": "jacobi3d_gcell_updates_per_s",
            "value": round(value, 3),
            "unit": "Gcell/s",
            "n_gpus": n,
            "steps": args.steps,
            "warmup": max(args.warmup, 3),
            "ms_per_step": round(secs / args.steps * 1e3, 5),
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "fp32",
            "data": "synthetic (field initialised to 0.5 with hot/cold Dirichlet spheres, as the reference app)",
            "impl": "own",
            "config": {
                "model": "jacobi3d 7-point radius-1",
                "global_batch": f"{X}x{X}x{X} cells",
                "seq_len": f"{args.size}^3 cells per GPU",
                "parallelism": f"3-D domain decomposition over {n} GPU(s), {deployment}, {mode} compute+halo-push kernel",
                "l2": "inputs larger than L2 (512 MB field per GPU vs 126 MB L2); no flush needed",
                "timing": "CUDA events per GPU around the K steps, max over GPUs and ranks",
            },
            "clocks": clocks,
            "e2e": {
                "value": round(e2e_value, 3),
                "unit": "Gcell/s",
                "ms_per_step": round(e2e_secs / args.steps * 1e3, 5),
                "h2d_bytes_per_step": 8 * nsub * (world if world > 1 else 1),
                "d2h_bytes_per_step": 4 * nsub * (world if world > 1 else 1),
                "results_valid": bool(finite),
                "timing": "host wall clock between barrier+sync points, max over ranks",
            },
            "gpu_launches": int(launches * (world if world > 1 else 1)),
            "exchange_us_per_iter": None,
            "checksum": checksum,
            "hbm_fraction_of_measured": None,
        }
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peak = json.load(f)["hbm_gbs"]
            out["hbm_fraction_of_measured"] = round((8.0 * float(args.size) ** 3 / (secs / args.steps)) / 1e9 / peak, 4)
        except Exception:
            pass
        print(json.dumps(out), flush=True)
    comm.finalize()


def run_reference(args):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import reference_arm
    reference_arm.main(args)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--size", type=int, default=512, help="cells per axis per GPU")
    ap.add_argument("--tile-z", type=int, default=16)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
