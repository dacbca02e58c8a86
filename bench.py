#!/usr/bin/env python
"""Headline benchmark: fused 2-D 5-point stencil, 16384 x 16384 fp64 cells per GPU.

Metric (BASELINE.json): 2D-stencil Gcell-updates/s at 1/2/4/8 B200, device
timed, max over ranks.  Weak scaling: every GPU owns a 16384^2 tile of a
periodic Py x Px Cartesian grid (1x1, 1x2, 2x2, 2x4); one step = one Jacobi
time step of every tile INCLUDING the halo exchange, which is fused into the same
kernel (peer stores over NVLink + device-side epoch flags).  By default the kernel is the
temporally blocked one (``--temporal 3``): every sweep over the grid advances THREE time steps
(ghost width 3, corners exchanged), bit-identical to three single steps, so the grid is read
and written once per three updates.

    python bench.py --gpus 1 --steps 20 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 \
        --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 3

Four arms run in the same process, on the same grid, for the same K time steps per timed
batch, all timed on the device (CUDA events bracketed by barrier + synchronize, max over ranks):

  1. ``value``                  fused stencil+halo kernel, T time steps per sweep (the product);
  2. ``one_time_step_per_sweep`` the same fused kernel with one exchange per time step (T = 1);
  3. ``unfused_baseline``       what the reference's loop does on this box
                                 (``stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172``:
                                 ``do { ExchangeData; Compute } while(...)`` with
                                 ``stencil2d/stencil2D.h:363-377``: all receives, all sends, one
                                 wait): pack -> NCCL ``batch_isend_irecv`` -> unpack -> this
                                 repo's stencil kernel on a plan WITHOUT neighbours, per time step;
  4. ``e2e``                    arm 1 through the public API with per-sweep pinned H2D of the
                                 coefficients and D2H + host read of the residual.

Every arm's K-step batch is repeated until ``--min-seconds`` of device time has passed; the
number reported is the MEDIAN batch (min / max alongside).  After the timed regions all three
implementations restart from the same field, run K steps, and their cores must be bit-identical
(``verified``).  ``vs_baseline`` = arm 1 / arm 3: the reference itself cannot be built here (no
MPI on the image, DESIGN.md "Reference arm"), so the host-driven equivalent measured in the same
run is the comparator.

Rank 0 prints ONE JSON line.  ``--impl reference`` reports ``unavailable`` (see above).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tile", type=int, default=16384, help="cells per side of each GPU's tile")
    ap.add_argument("--radius", type=int, default=1)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--temporal", type=int, default=0, choices=[0, 1, 2, 3, 4],
                    help="time steps per kernel sweep: 2-4 = temporal blocking (radius 1 only), "
                         "1 = one step per sweep, 0 = 3 where possible")
    ap.add_argument("--dims", default="", help="process grid PYxPX (default: balanced, like MPI_Dims_create)")
    ap.add_argument("--th", type=int, default=0, help="explicit per-GPU tile height (experiments)")
    ap.add_argument("--tw", type=int, default=0, help="explicit per-GPU tile width (experiments)")
    ap.add_argument("--min-seconds", type=float, default=2.0,
                    help="repeat the K-step batch of the headline arm until this much device time "
                         "has passed (the other arms get half of it)")
    ap.add_argument("--no-baseline", action="store_true", help="skip arms 2 and 3 (experiments)")
    return ap.parse_args(argv)


def reference_arm(args) -> int:
    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        print(json.dumps({
            "impl": "reference",
            "unavailable": "reference is a set of C++/MPI main() programs with no setup.py/"
                           "pyproject (pip: 'not installable') and this image has no MPI "
                           "(mpicxx, mpi.h, libmpi, mpi4py all absent), so it cannot be built",
        }))
    return 0


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL writes its
    version line to stdout when NCCL_DEBUG is set on the box), so file descriptor 1 is pointed at
    stderr for the duration of the run and the JSON line goes to the original stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(real, "w")


def _stats(ms_list, steps):
    per = sorted(m / steps for m in ms_list)
    return {"median": statistics.median(per), "min": per[0], "max": per[-1], "batches": len(per)}


def main(argv=None) -> int:
    args = parse_args(argv)
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import cuda_mpi_scratch_b200 as cms
    from cuda_mpi_scratch_b200.models.stencil2d import Stencil2D, jacobi_coeffs
    from cuda_mpi_scratch_b200.ops.unfused_halo import UnfusedHalo
    from cuda_mpi_scratch_b200.utils.timing import ClockSampler

    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device"}))
        return 1
    result_out = _claim_stdout()
    world = cms.init()
    n = world.size
    if n != args.gpus:
        if world.rank == 0:
            print(f"warning: --gpus {args.gpus} but WORLD_SIZE={n}; using {n}", file=sys.stderr)
    dims = cms.dims_create(n, 2)
    if args.scaling == "strong" and not args.dims and args.tile % n == 0 and args.tile // n >= 512:
        # Fixed total work: row slabs keep every tile as wide as the grid.  The strip kernel wants
        # wide tiles (11 full-width strips instead of 3, profiles/r2/tile_shapes_1gpu.txt) and the
        # whole halo is then contiguous rows; the balanced 2 x 4 grid is --dims 2x4.
        dims = (n, 1)
    if args.dims:
        dims = tuple(int(v) for v in args.dims.lower().split("x"))
        if len(dims) != 2 or dims[0] * dims[1] != n:
            raise SystemExit(f"--dims {args.dims} does not describe {n} ranks")
    if args.scaling == "weak":
        th = tw = args.tile
    else:  # strong: total work fixed at tile^2, split over the grid
        th, tw = args.tile // dims[0], args.tile // dims[1]
    if args.th and args.tw:
        th, tw = args.th, args.tw
    R = args.radius
    T = args.temporal or (3 if R == 1 else 1)
    K, Wm = args.steps, max(args.warmup, 3)
    coeffs = jacobi_coeffs(R)
    dev = world.device
    device = f"cuda:{dev}"
    stream = torch.cuda.current_stream(dev)
    cart = cms.CartTopology.create(dims, True)
    cr, cc = cart.coords(world.rank)

    # synthetic data: smooth deterministic field generated per tile on the device (state, like
    # weights: it stays resident; the per-step INPUTS of the e2e arm are the coefficients)
    ys = torch.arange(th, device=device, dtype=torch.float64) + cr * th
    xs = torch.arange(tw, device=device, dtype=torch.float64) + cc * tw

    def init_field(p, collective=True):
        p.fill(0.0, 0.0)
        p.core.copy_(torch.sin(ys * 1e-3)[:, None] * torch.cos(xs * 2e-3)[None, :])
        p.set_coeffs(coeffs)
        torch.cuda.synchronize(dev)
        if collective:
            p.exchange()
            torch.cuda.synchronize(dev)
            world.barrier()

    def make_plan(temporal):
        p = Stencil2D.create(world, (th, tw), radius=R, kind="star", dims=dims, periods=True,
                             temporal=temporal)
        init_field(p)
        return p

    def timed_batches(run_batch, min_s, max_batches=2000, wall=False):
        """K-step batches, each bracketed by barrier + synchronize on both sides and timed with
        CUDA events on the launching stream (``wall``: host clock instead, for the e2e arm);
        repeated until ``min_s`` of timed work; per-batch MAX over ranks."""
        def one():
            world.barrier()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(stream)
            run_batch()
            e1.record(stream)
            torch.cuda.synchronize(dev)
            t1 = time.perf_counter()
            return (t1 - t0) * 1e3 if wall else e0.elapsed_time(e1)
        first = world.allreduce_max(one())
        reps = int(min(max_batches, max(3, -(-min_s * 1e3 // max(first, 1e-3)))))
        local = [one() for _ in range(reps)]
        world.barrier()
        return world.allreduce_max_vec(local)

    # ---------------- arm 1: fused, T time steps per sweep, device timed ----------------
    st = make_plan(T)
    sampler = ClockSampler(dev).start() if world.rank == 0 else None
    if sampler is not None:
        sampler.wait_first(2.0)
    st.step(Wm)
    st.step(K)                                   # one untimed batch of the timed shape
    torch.cuda.synchronize(dev)
    launches0 = st.kernel_launches
    st.step(K)
    launches = st.kernel_launches - launches0    # kernels of ONE K-step batch
    torch.cuda.synchronize(dev)
    if sampler is not None:
        sampler.window_begin()
    ms_main = timed_batches(lambda: st.step(K), args.min_seconds)
    if sampler is not None:
        sampler.window_end()
    st.check()
    main_stats = _stats(ms_main, K)

    # ---------------- arm 4: end to end through the public API ----------------
    # every sweep: H2D of that sweep's coefficients from pinned memory, one fused sweep with
    # residual, D2H of the residual into pinned memory, and the host READS it -- one sweep
    # late, so that the next sweep is already queued while the copy of this one completes
    # (two pinned slots per direction inside Stencil2D).
    coef_host = torch.tensor(coeffs, dtype=torch.float64).pin_memory()
    h2d, d2h = coef_host.numel() * 8, 8
    chunks = [T] * (K // T) + ([K % T] if K % T else [])
    last_res = [0.0]

    def e2e_batch():
        pend = None
        for i, nsub in enumerate(chunks):        # exactly K time steps, one sweep per iteration
            # Stencil2D.step_io: H2D of this sweep's coefficients (pinned) and D2H of its residual
            # on the plan's copy streams; ghost cells of the user-visible tile are only needed
            # when the batch ends
            t = st.step_io(nsub, coef_host, residual=True, drain=(i == len(chunks) - 1))
            if pend is not None:
                last_res[0] = st.wait_io(pend)   # blocks until the PREVIOUS sweep's residual is on the host
            pend = t
        last_res[0] = st.wait_io(pend)

    for _ in range(2):
        e2e_batch()
    if sampler is not None:
        sampler.window_begin()
    ms_e2e = timed_batches(e2e_batch, args.min_seconds / 2, wall=True)
    if sampler is not None:
        sampler.window_end()
    clocks = sampler.stop() if sampler is not None else None
    st.check()
    e2e_stats = _stats(ms_e2e, K)

    # ---------------- arm 2: fused, one time step per sweep ----------------
    single_stats = st1 = None
    if T > 1 and not args.no_baseline:
        st1 = make_plan(1)
        st1.step(Wm)
        torch.cuda.synchronize(dev)
        single_stats = _stats(timed_batches(lambda: st1.step(K), args.min_seconds / 2), K)
        st1.check()

    # ---------------- arm 3: UNFUSED baseline (pack, NCCL send/recv, unpack, separate kernel) ---
    unf_stats = loc = None
    if R == 1 and not args.no_baseline:
        grp = Stencil2D.create_group(cms.LocalGroup(1, devices=[dev]), (th, tw), 1, "star",
                                     dims=(1, 1), periods=False)
        loc = grp.ranks[0]                        # a plan whose 8 neighbours are absent never
        init_field(loc, collective=False)         # pushes, waits or signals inside the kernel
        halo = UnfusedHalo(cart, world.rank, (th, tw), 1, device)

        def unfused_steps(k=K):
            for _ in range(k):
                halo.exchange(loc.core, loc.tile)     # views of the CURRENT buffer
                loc.step(1, drain=False)

        unfused_steps(Wm)
        torch.cuda.synchronize(dev)
        unf_stats = _stats(timed_batches(unfused_steps, args.min_seconds / 2), K)
        loc.check()

    # ---------------- verification (outside every timed region) ----------------
    # all implementations restart from the same field and advance K time steps: the cores must
    # agree BIT FOR BIT (temporal blocking vs single steps vs host-driven exchange)
    verified = None
    if not args.no_baseline and (st1 is not None or loc is not None):
        init_field(st)
        st.step(K)
        torch.cuda.synchronize(dev)
        bad = 0.0
        if st1 is not None:
            init_field(st1)
            st1.step(K)
            torch.cuda.synchronize(dev)
            bad += 0.0 if torch.equal(st.core, st1.core) else 1.0
        if loc is not None:
            init_field(loc, collective=False)
            unfused_steps(K)
            torch.cuda.synchronize(dev)
            bad += 0.0 if torch.equal(st.core, loc.core) else 1.0
        verified = world.allreduce_sum(bad) == 0.0

    cells = float(th) * float(tw) * n
    rate = lambda ms_per_step: cells / (ms_per_step * 1e-3) / 1e9
    value = rate(main_stats["median"])
    if world.rank == 0:
        bytes_per_cell = 16.0 / T   # one read + one write of the grid per sweep of T updates
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        per_gpu = value / n
        nsweeps = len(chunks)

        def arm(stats, what):
            return None if stats is None else {
                "value": round(rate(stats["median"]), 3), "unit": "Gcell/s",
                "ms_per_step": round(stats["median"], 5),
                "ms_per_step_min": round(stats["min"], 5), "ms_per_step_max": round(stats["max"], 5),
                "batches": stats["batches"], "what": what}

        unf = arm(unf_stats, "UNFUSED baseline, same grid and K: per time step pack (torch copies) -> "
                             "NCCL batch_isend_irecv of the 4 edges -> unpack -> this repo's stencil "
                             "kernel on a plan without neighbours; device timed, median batch")
        one = arm(single_stats, "fused kernel, one time step (one halo exchange) per sweep "
                                "(--temporal 1); device timed, median batch")
        out = {
            "metric": "2D-stencil Gcell-updates/s (5-point fp64, fused halo exchange)",
            "value": round(value, 3), "unit": "Gcell/s", "n_gpus": n, "steps": K,
            "warmup": args.warmup, "ms_per_step": round(main_stats["median"], 5),
            "ms_per_step_min": round(main_stats["min"], 5),
            "ms_per_step_max": round(main_stats["max"], 5),
            "timed_batches": main_stats["batches"],
            "timed_region_s": round(sum(ms_main) * 1e-3, 3),
            "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None if unf is None else round(value / unf["value"], 3),
            "vs_baseline_what": "value / unfused_baseline.value, both measured in this run (the "
                                "reference publishes no number and cannot be built here)",
            "dtype": "fp64", "data": "synthetic",
            "impl": "ours",
            "verified": verified,
            "verified_what": f"after K={K} steps from the same field the cores of the fused T={T} run, "
                             "the fused T=1 run and the unfused baseline are bit-identical on every rank",
            "config": {"model": f"stencil2d star radius {R} ({4 * R + 1}-point) Jacobi, periodic"
                                + (f", {T} time steps per kernel sweep (temporal blocking, bit-identical "
                                   "to single steps)" if T > 1 else ""),
                       "tile_per_gpu": [th, tw], "global_grid": [th * dims[0], tw * dims[1]],
                       "process_grid": list(dims), "global_batch": int(cells), "seq_len": None,
                       "parallelism": f"cart{dims[0]}x{dims[1]} 2-D domain decomposition",
                       "l2": "inputs larger than L2 (2 x %.2f GiB per GPU vs 126 MB)"
                             % (th * tw * 8 / 2 ** 30),
                       "halo": "fused: peer stores + epoch flags inside the stencil kernel",
                       "time_steps_per_kernel_sweep": T,
                       "kernel": ("stencil_tstrip_kernel (strips marched top to bottom, time levels in registers)"
                                  if st.info().threads == 384 else "stencil_fused_kernel (tiles)"),
                       "timing": "K-step batches, each between barrier+synchronize, CUDA events, max "
                                 "over ranks per batch, repeated for >= %.1f s; value = median batch"
                                 % args.min_seconds},
            "roofline": {"hbm_gbs_measured": hbm,
                         "achieved_gbs_per_gpu": round(per_gpu * bytes_per_cell, 1),
                         "fraction_of_measured_hbm": round(per_gpu * bytes_per_cell / hbm, 4)},
            "clocks": clocks,
            "unfused_baseline": unf,
            "one_time_step_per_sweep": one,
            "fused_vs_unfused_one_step_per_sweep":
                None if (unf is None or one is None) else round(one["value"] / unf["value"], 3),
            "e2e": {"value": round(rate(e2e_stats["median"]), 3), "unit": "Gcell/s",
                    "ms_per_step": round(e2e_stats["median"], 5),
                    "ms_per_step_min": round(e2e_stats["min"], 5),
                    "ms_per_step_max": round(e2e_stats["max"], 5), "batches": e2e_stats["batches"],
                    "h2d_bytes_per_step": round(h2d * nsweeps / K, 3),
                    "d2h_bytes_per_step": round(d2h * nsweeps / K, 3),
                    "vs_unfused_baseline": None if unf is None else
                        round(rate(e2e_stats["median"]) / unf["value"], 3),
                    "what": f"Stencil2D.step_io({T}, coeffs, residual=True) + wait_io(previous call) "
                            "every kernel sweep: H2D of the sweep's coefficients from pinned memory "
                            "and D2H of its residual into pinned memory on the plan's copy streams, "
                            "host reads the residual one sweep late (next sweep already queued); "
                            "wall clock per K-step batch, max over ranks, median batch",
                    "last_residual_sumsq": last_res[0]},
            "gpu_launches": launches * main_stats["batches"],
            "gpu_launches_per_batch": launches,
            "launches_note": "kernels of this repo launched inside the timed region of the headline "
                             "arm = kernels of one K-step batch (counted by the native plan) x timed "
                             "batches.  Per batch: 1 fused stencil+halo kernel per sweep "
                             f"({T} time step(s) each; its last CTA advances the device step counter), "
                             "then 1 flag drain and 1 ghost-column unpack",
        }
        print(json.dumps(out), file=result_out, flush=True)
    world.barrier()
    for p in (st, st1, loc):
        if p is not None:
            p.close()
    world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
