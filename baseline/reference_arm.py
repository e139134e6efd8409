"""bench.py --impl reference: run the unmodified reference's own binaries (baseline/_ref/bin) on the same metric and
config as the product arm and print the same JSON line with "impl": "reference".

  value                  baseline/_ref/bin/jacobi3d 512 512 512 -n <steps+warmup>   (it weak-scales the size itself)
  exchange_us_per_iter   baseline/_ref/bin/weak 512 512 512 <steps> --peer --kernel  (radius 3, 4 x fp32, 26 directions)
  astaroth_gcell_per_s   none: the reference's astaroth-sim runs one fp32 field with stand-in compute and prints no timing

Deployment: the reference as 1 rank driving N GPUs (its own default mode: a rank takes every visible device), because
this image has no MPI; under torchrun only rank 0 runs it and the other ranks idle. The reference reports per-iteration
HOST wall time (MPI_Wtime around launch..cudaStreamSynchronize, bin/jacobi3d.cu:267-341) as min and trimean over all
iterations including its first ones; the metric uses the trimean. Its exchange time is dd.timeExchange_ / iterations
(host wall clock with a barrier per call, bin/weak.cu:186-193)."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def unavailable(why):
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def _run(cmd, env, sampler_cls):
    """Run one reference binary; clocks are sampled only while the GPU is busy (the binary's set-up phase, during which
    the GPU idles at its lowest clock, is not part of any timed region)."""
    sampler = None
    try:
        sampler = sampler_cls(0, period_s=0.005, busy_only=True).start() if sampler_cls else None
    except Exception:  # noqa: BLE001
        sampler = None
    try:
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=3000, env=env, cwd="/tmp")
    finally:
        clocks = sampler.stop() if sampler else None
    return r, clocks


def main(args, config, metric):
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
    try:
        sys.path.insert(0, os.path.dirname(HERE))
        from stencil2_b200.utils.clocks import ClockSampler
    except Exception:  # noqa: BLE001
        ClockSampler = None
    warm = max(args.warmup, 3)
    iters = args.steps + warm
    try:
        r, clocks = _run([exe, str(args.size), str(args.size), str(args.size), "-n", str(iters)], env, ClockSampler)
    except Exception as e:  # noqa: BLE001
        unavailable(f"reference jacobi3d failed to run: {e}")
        return
    line = [l for l in r.stdout.splitlines() if l.startswith("jacobi3d,")]
    if r.returncode != 0 or not line:
        unavailable("reference jacobi3d exited %d: %s" % (r.returncode, (r.stderr or r.stdout)[-300:].replace("\n", " | ")))
        return
    f = line[-1].split(",")
    methods, ranks, dev = f[1], int(f[2]), int(f[3])
    x, y, z = int(f[4]), int(f[5]), int(f[6])
    tmin, ttri = float(f[7]), float(f[8])
    cells = float(x) * y * z

    exchange = None
    if not getattr(args, "skip_exchange", False):
        wexe = os.path.join(HERE, "_ref", "bin", "weak")
        try:
            rw, _ = _run([wexe, str(args.size), str(args.size), str(args.size), str(args.steps), "--peer", "--kernel"], env, None)
            wl = [l for l in rw.stdout.splitlines() if l.startswith("weak,")]
            if rw.returncode == 0 and wl:
                w = wl[-1].split(",")
                ex_s = float(w[21]) / int(w[10])
                exchange = {"us_per_exchange_host": round(ex_s * 1e6, 3), "bytes_per_exchange": int(w[8]) + int(w[9]),
                            "config": f"radius 3, 4 x fp32, 26 directions, {w[2]}x{w[3]}x{w[4]} cells, methods {w[1]}",
                            "timing": "the reference's own dd.timeExchange_ / iterations (host wall clock, barrier per call)"}
        except Exception:  # noqa: BLE001
            exchange = None

    print(json.dumps({
        "metric": metric, "value": round(cells / ttri / 1e9, 3), "unit": "Gcell/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": warm, "ms_per_step": round(ttri * 1e3, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic (reference app's own initialisation)",
        "impl": "reference",
        "config": config,
        "deployment": f"reference jacobi3d, {ranks} rank x {dev} GPU(s), methods {methods}, overlap on",
        "timing": "the reference's own per-iteration host wall clock (trimean over all %d iterations; min %.6f s)" % (iters, tmin),
        "build": "unmodified sources, nvcc sm_100 -O3, single-rank MPI shim (no MPI on this image), nvtx3 headers",
        "e2e": {"value": round(cells / ttri / 1e9, 3), "unit": "Gcell/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "the reference app has no per-step host I/O; its timing is already host wall clock"},
        "clocks": clocks,
        "gpu_launches": None,
        "exchange_us_per_iter": exchange["us_per_exchange_host"] if exchange else None,
        "exchange": exchange,
        "astaroth_gcell_per_s": None,
        "astaroth": {"note": "the reference's astaroth-sim runs one fp32 field with stand-in compute and prints no timing"},
    }), flush=True)
