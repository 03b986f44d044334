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


def bench(a, tau: float):
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
    nrep = max(2, min(3, a.repeats))  # whole-program runs; the first one also warms driver + page cache
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
        flag = [err]
        gathered = [None] * ws
        dist.all_gather_object(gathered, err)
        err = next((g for g in gathered if g), None)
        dist.destroy_process_group()
    if err:
        emit({"impl": "reference", "unavailable": err.replace("\n", " ")[:400]})
        return
    if rank == 0:
        t_ms = min(times[1:]) if len(times) > 1 else times[0]
        cells = float(N + 1) ** 3
        ea, er, extras = last
        emit({
            "metric": "Gcell-updates/s (device-timed, max over ranks)", "value": cells * K / (t_ms * 1e-3) / 1e9,
            "unit": "Gcell-updates/s", "n_gpus": a.gpus, "steps": K, "warmup": a.warmup, "ms_per_step": t_ms / K,
            "higher_is_better": True, "scaling": getattr(a, "scaling", "strong"), "vs_baseline": None, "dtype": "fp64",
            "data": "synthetic (analytic initial condition)", "impl": "reference",
            "config": {"model": f"cuda_sol N={N} L={L} tau={tau} timesteps={K}", "global_batch": int(cells), "seq_len": K,
                       "parallelism": f"{ws} MPI ranks (MPI_Dims_create grid), 1 rank per GPU",
                       "transport": "MPI stand-in (baseline/mpi_shim/mpi.h: shared-memory mailboxes); host-staged faces "
                                    "as in the reference",
                       "timer": "reference's own rank-0 cudaEvent pair ('numerical solution calculated in'); best of "
                                f"{len(times) - 1 if len(times) > 1 else 1} whole-program runs after one warm-up run",
                       "solve_ms_all": times, "shim_wall_ms_max_over_ranks": walls,
                       "reference_timers_ms": extras,
                       "max_abs_error_last_layer": ea[-1] if ea else None,
                       "max_rel_error_last_layer": er[-1] if er else None},
            "gpu_launches": None,
        })
    shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    sys.exit(0)
