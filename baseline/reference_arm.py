"""bench.py --impl reference: run the unmodified reference's own jacobi3d binary (baseline/_ref/bin/jacobi3d) on the
same metric and config as the product arm and print the same JSON line with "impl": "reference".

Deployment: the reference as 1 rank driving N GPUs (its own default mode: a rank takes every visible device), because
this image has no MPI; under torchrun only rank 0 runs it and the other ranks idle. The reference reports per-iteration
HOST wall time (MPI_Wtime around launch..cudaStreamSynchronize, bin/jacobi3d.cu:267-341) as min and trimean over all
iterations including its first ones; the metric uses the trimean. It weak-scales the positional size itself."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def unavailable(why):
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def main(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    exe = os.path.join(HERE, "_ref", "bin", "jacobi3d")
    if not os.path.exists(exe):
        try:
            sys.path.insert(0, HERE)
            import build_reference
            build_reference.build()
        except Exception as e:  # noqa: BLE001
            unavailable(f"reference not built and cannot be built here: {e}")
            return
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(args.gpus))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    iters = args.steps + max(args.warmup, 3)
    cmd = [exe, str(args.size), str(args.size), str(args.size), "-n", str(iters)]
    sampler = None
    try:  # clocks are sampled over the whole run of the binary (set-up included; the app exposes no hook for less)
        sys.path.insert(0, os.path.dirname(HERE))
        from stencil2_b200.utils.clocks import ClockSampler
        sampler = ClockSampler(0, period_s=0.02).start()
    except Exception:  # noqa: BLE001
        sampler = None
    try:
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=3000, env=env, cwd="/tmp")
    except Exception as e:  # noqa: BLE001
        unavailable(f"reference jacobi3d failed to run: {e}")
        return
    finally:
        clocks = sampler.stop() if sampler else None
    line = [l for l in r.stdout.splitlines() if l.startswith("jacobi3d,")]
    if r.returncode != 0 or not line:
        unavailable("reference jacobi3d exited %d: %s" % (r.returncode, (r.stderr or r.stdout)[-300:].replace("\n", " | ")))
        return
    f = line[-1].split(",")
    methods, ranks, dev = f[1], int(f[2]), int(f[3])
    x, y, z = int(f[4]), int(f[5]), int(f[6])
    tmin, ttri = float(f[7]), float(f[8])
    cells = float(x) * y * z
    print(json.dumps({
        "metric": "jacobi3d_gcell_updates_per_s", "value": round(cells / ttri / 1e9, 3), "unit": "Gcell/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ttri * 1e3, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic (reference app's own initialisation)",
        "impl": "reference",
        "config": {"model": "jacobi3d 7-point radius-1", "global_batch": f"{x}x{y}x{z} cells", "seq_len": f"{args.size}^3 cells per GPU",
                   "parallelism": f"reference jacobi3d, {ranks} rank x {dev} GPU(s), methods {methods}, overlap on",
                   "timing": "the reference's own per-iteration host wall clock (trimean over all %d iterations; min %.6f s)" % (iters, tmin),
                   "build": "unmodified sources, nvcc sm_100 -O3, single-rank MPI shim (no MPI on this image), nvtx3 headers"},
        "e2e": {"value": round(cells / ttri / 1e9, 3), "unit": "Gcell/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "the reference app has no per-step host I/O; its timing is already host wall clock"},
        "clocks": clocks,
        "gpu_launches": None,
    }), flush=True)
