"""Run the UNMODIFIED reference programs (baseline/_ref, built by build_reference.sh) for bench.py
--impl reference and for the parity tests.

The reference needs MPI; the image has none, so the binaries are linked against the single-node
stand-in baseline/mpi_shim/mpi.h (processes + shared-memory mailboxes). Under torchrun every
rank of the launcher becomes one MPI rank of the reference (rank r -> GPU r % ndev, as the reference
does itself, ref: cuda_sol.cpp:519). The number reported is the reference's own
"numerical solution calculated in ... ms" (rank 0, cudaEvent pair, ref: cuda_sol.cpp:382,422-428);
the stand-in additionally records each rank's wall time over the same region (first MPI_Barrier ->
first MPI_Reduce) so that a max over ranks is available without touching the reference sources.
"""
from __future__ import annotations

import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import uuid

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
BIN = os.path.join(REF, "bin")


def available(prog: str = "cuda_sol") -> bool:
    return os.path.exists(os.path.join(BIN, prog))


def parse_output(path: str):
    """-> (solve_ms, max_abs[], max_rel[], extras) from a reference output file."""
    txt = open(path).read()
    m = re.search(r"numerical solution calculated in ([0-9.eE+-]+)\s*ms", txt)
    ms = float(m.group(1)) if m else float("nan")
    ea, er = [], []
    for mm in re.finditer(r"max abs and rel errors on layer (\d+): (\S+) (\S+)", txt):
        ea.append(float(mm.group(2)))
        er.append(float(mm.group(3)))
    extras = {k: float(v) for k, v in re.findall(r"total ([A-Za-z\- ]+ time): ([0-9.eE+-]+)", txt)}
    return ms, ea, er, extras


def run_program(prog: str, argv, nranks: int = 1, rank: int = 0, job: str | None = None, cwd: str | None = None,
                binary: str | None = None, timeout: float = 900.0, extra_env=None):
    """Run one rank of a reference program (all ranks must be started, each with its own `rank`)."""
    exe = binary or os.path.join(BIN, prog)
    env = dict(os.environ)
    env.update({"SHIM_RANK": str(rank), "SHIM_SIZE": str(nranks), "SHIM_JOB": job or uuid.uuid4().hex[:12]})
    if cwd:
        env["SHIM_TIMING_DIR"] = cwd
    if extra_env:
        env.update(extra_env)
    return subprocess.run([exe] + [str(a) for a in argv], cwd=cwd, env=env, capture_output=True, text=True,
                          timeout=timeout)


def run_all_ranks_local(prog: str, argv, nranks: int, cwd: str, binary: str | None = None, timeout: float = 900.0,
                        extra_env=None):
    """Start all ranks from this process (tests / single-launcher use)."""
    job = uuid.uuid4().hex[:12]
    exe = binary or os.path.join(BIN, prog)
    procs = []
    for r in range(nranks):
        env = dict(os.environ)
        env.update({"SHIM_RANK": str(r), "SHIM_SIZE": str(nranks), "SHIM_JOB": job, "SHIM_TIMING_DIR": cwd})
        if extra_env:
            env.update(extra_env)
        procs.append(subprocess.Popen([exe] + [str(a) for a in argv], cwd=cwd, env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, o, e))
    return outs


def dims_create(nranks: int):
    """Process grid the reference gets from MPI_Dims_create (ref: cuda_sol.cpp:478): balanced, non-increasing.
    Plain Python on purpose — this arm loads nothing of the repo's native code."""
    primes, m, f = [], nranks, 2
    while f * f <= m:
        while m % f == 0:
            primes.append(f)
            m //= f
        f += 1
    if m > 1:
        primes.append(m)
    d = [1, 1, 1]
    for f in sorted(primes, reverse=True):
        d[d.index(min(d))] *= f
    return tuple(sorted(d, reverse=True))


def staged_bytes_per_step(N: int, dims):
    """Bytes the reference moves between host and device inside its timed region, per step and per rank
    (rank 0's block): exchange() copies all six padded send faces device->host and all six receive faces
    host->device whether or not a neighbour exists (ref: cuda_sol.cpp:236-243, 296-304); calculate_errors
    copies two arrays of one double per 512-cell block device->host (ref: cuda_sol_kernels.cu:80-88)."""
    X, Y, Z = ((N + 1) // d for d in dims)
    faces = 2 * ((Y + 2) * (Z + 2) + (X + 2) * (Z + 2) + (X + 2) * (Y + 2)) * 8
    blocks = -(-((X + 2) * (Y + 2) * (Z + 2)) // 512)
    return faces, faces + 2 * blocks * 8


def bench(a, tau: float, problem_config):
    """bench.py --impl reference: same metric/config as our arm, one JSON line from rank 0."""
    rank = int(os.environ.get("RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))

    def emit(d):
        if rank == 0:
            print(json.dumps(d), flush=True)

    if not available("cuda_sol"):
        emit({"impl": "reference", "unavailable": "baseline/_ref/bin/cuda_sol is not built (run baseline/build_reference.sh "
              "on a box that has /root/reference)"})
        return
    N, K = a.n, a.steps
    if getattr(a, "scaling", "strong") == "weak":
        N = int(round((a.n + 1) * (a.gpus ** (1.0 / 3.0)))) - 1
    L = a.length
    T = K * tau
    # agree on a job id + scratch directory across the launcher's ranks
    if ws > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=ws)
        box = [uuid.uuid4().hex[:12] if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        job = box[0]
    else:
        job = uuid.uuid4().hex[:12]
    work = os.path.join(tempfile.gettempdir(), f"w3d_ref_{job}")
    os.makedirs(work, exist_ok=True)
    argv = [N, 1, L, L, L, repr(T), K]
    face = (N + 3) * (N + 3) * 8 + 4096
    times, walls, last = [], [], None
    clk = None
    if rank == 0:
        try:  # SM clocks and throttle reasons while the reference runs (plain-Python NVML sampler, loaded by
            # path so that nothing else of the repo's package is imported into this arm)
            import importlib.util
            spec = importlib.util.spec_from_file_location(
                "w3d_clocks", os.path.join(os.path.dirname(HERE), "3d-wave-equation-mpi-cuda_b200", "utils", "clocks.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            clk = mod.ClockSampler(0, 0.02)
            clk.__enter__()
        except Exception:  # noqa: BLE001
            clk = None
    # whole-program runs: one untimed run (driver, page cache, the W warm-up steps) and then the timed ones;
    # the median is reported, the same statistic as our arm
    nrep = 1 + max(1, min(3, a.repeats))
    err = None
    for rep in range(nrep):
        r = run_program("cuda_sol", argv, nranks=ws, rank=rank, job=f"{job}_{rep}", cwd=work,
                        extra_env={"SHIM_CAP_BYTES": str(face)})
        if r.returncode != 0:
            err = f"cuda_sol exited with {r.returncode}: {(r.stderr or r.stdout)[-300:]}"
        if ws > 1:
            import torch.distributed as dist
            dist.barrier()
        if err:
            break
        if rank == 0:
            import glob
            outs = glob.glob(os.path.join(work, f"output_N{N}_Np{ws}_Ng*_cuda.txt"))
            ms, ea, er, extras = parse_output(outs[0])
            w = []
            for q in range(ws):
                try:
                    w.append(float(open(os.path.join(work, f"rank_{q}.txt")).read()))
                except Exception:  # noqa: BLE001
                    pass
            times.append(ms)
            walls.append(max(w) if w else float("nan"))
            last = (ea, er, extras)
        if ws > 1:
            import torch.distributed as dist
            dist.barrier()
    if ws > 1:
        import torch.distributed as dist
        gathered = [None] * ws
        dist.all_gather_object(gathered, err)
        err = next((g for g in gathered if g), None)
        dist.destroy_process_group()
    clocks = {}
    if clk is not None:
        clk.__exit__(None, None, None)
        clocks = clk.summary()
    if err:
        emit({"impl": "reference", "unavailable": err.replace("\n", " ")[:400]})
        return
    if rank == 0:
        import statistics
        timed = times[1:] if len(times) > 1 else times
        t_ms = statistics.median(timed)
        cells = float(N + 1) ** 3
        ea, er, extras = last
        dims = dims_create(ws)
        out = {
            "metric": "Gcell-updates/s (device-timed, max over ranks)", "value": cells * K / (t_ms * 1e-3) / 1e9,
            "unit": "Gcell-updates/s", "n_gpus": a.gpus, "steps": K, "warmup": a.warmup, "ms_per_step": t_ms / K,
            "higher_is_better": True, "scaling": getattr(a, "scaling", "strong"), "vs_baseline": None, "dtype": "fp64",
            "data": "synthetic (analytic initial condition)", "impl": "reference",
            "config": problem_config(N, L, tau, K, dims, cells),
            "detail": {"program": "baseline/_ref/bin/cuda_sol (unmodified reference sources, ARCH=sm_100a)",
                       "ranks": f"{ws} MPI ranks (MPI_Dims_create grid), 1 rank per GPU",
                       "transport": "MPI stand-in (baseline/mpi_shim/mpi.h: shared-memory mailboxes); host-staged faces "
                                    "as in the reference",
                       "timer": "reference's own rank-0 cudaEvent pair ('numerical solution calculated in'); median of "
                                f"{len(timed)} whole-program runs after one warm-up run",
                       "repeats": len(timed), "solve_ms_all": times, "shim_wall_ms_max_over_ranks": walls,
                       "reference_timers_ms": extras,
                       "max_abs_error_last_layer": ea[-1] if ea else None,
                       "max_rel_error_last_layer": er[-1] if er else None},
            "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                       "reasons": clocks.get("reasons", []), "samples": clocks.get("samples"),
                       "note": "sampled over the whole-program runs (process start-up and allocation included)"},
            "gpu_launches": None,
        }
        # End to end: the reference's timed region IS its end-to-end path — every step stages the faces and the
        # block maxima through pinned host memory and syncs the host (ref: cuda_sol.cpp:382-428). The MPI
        # stand-in clocks the same region on every rank (first MPI_Barrier -> first MPI_Reduce); max over ranks.
        wt = [w for w in (walls[1:] if len(walls) > 1 else walls) if w == w and w > 0]
        if wt:
            w_ms = statistics.median(wt)
            h2d, d2h = staged_bytes_per_step(N, dims)
            out["e2e"] = {"value": cells * K / (w_ms * 1e-3) / 1e9, "unit": "Gcell-updates/s", "ms_per_step": w_ms / K,
                          "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                          "note": "wall clock of the reference's whole solve (first MPI_Barrier -> first MPI_Reduce), max "
                                  "over ranks, measured by the MPI stand-in; per step it copies six padded faces each way "
                                  "through pinned host memory plus two block-maxima arrays device->host; bytes are per rank"}
        emit(out)
    shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    sys.exit(0)
