#!/usr/bin/env python
"""bench.py — instance-timepoints/sec of the batched dbmixer.sp transient (BASELINE.json metric).

A "step" = one pass of the hot path over one batch: DC operating point + the full
`.TRAN 1e-13 5e-9` (50,000 backward-Euler steps) for every instance of the rank's shard
(runTransientAnalysisBackwardEuler semantics, src/tanalisis.cpp:83-424; the DC solve is inside
the reference's transient driver, so it is inside the timed region here too).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
  python bench.py --impl reference ...                      the reference's own CPU path (oracle/_ref)

Workload at N GPUs: 8,192 instances per GPU (65,536 at N=8 — the configuration the north-star target
is quoted on), Monte-Carlo parameters (5% uniform, PCG64 seed 20261104), shards of ONE global batch,
no data-path collective ("weak" scaling).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 20261104
METRIC = "instance-timepoints/sec, batched dbmixer.sp transient"
UNIT = "instance-timepoints/s"


def load_netlist_text(name):
    """The reference's input netlists (tests/buffer.sp, tests/dbmixer.sp) travel inside the golden fixtures."""
    g = np.load(os.path.join(ROOT, "tests", "golden", f"{name}_nominal.npz"))
    return str(g["netlist"])


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--netlist", default="dbmixer", choices=["dbmixer", "buffer"])
    ap.add_argument("--instances-per-gpu", type=int, default=8192)
    ap.add_argument("--tran-steps", type=int, default=0, help="debug only: truncate the transient (0 = full .TRAN)")
    ap.add_argument("--out-stride", type=int, default=1)
    ap.add_argument("--stamp-instances", type=int, default=1 << 20)
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances in the CPU baseline sample (0 = 2 per thread)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--strong-total", type=int, default=65536,
                    help="strong-scaling record: this many instances split over the N GPUs, one timed run (0 = skip)")
    ap.add_argument("--strong-out-stride", type=int, default=100)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------------------------
def cpu_reference_leg(nl_name, n_inst, n_threads, tran_steps):
    """The UNMODIFIED reference (oracle/_ref), one instance per host thread at a time.
    Nothing of the product is on this path: the netlist is parsed by the reference's own parser (O.Ref), the
    Monte-Carlo draw is the oracle's restatement of the recipe, the timed wall is inside ref_run_batch."""
    from oracle import oraclelib as O
    ref = O.Ref(load_netlist_text(nl_name))
    topo = ref.topology()
    tran = ref.tran_cfg()
    ref.close()
    P = O.monte_carlo(topo, n_inst, SEED)  # [n_inst, P]
    if tran_steps:
        tran["tstop"] = tran["tstep"] * tran_steps
    texts = [O.netlist_from_params(topo, P[b], tran=tran) for b in range(n_inst)]
    wall, pts = O.ref_run_batch(texts, n_threads)
    return wall, pts


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oraclelib as O
    if not O.ref_available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libcsim_ref.so not present"}))
        return
    cores = O.ref_hardware_threads()
    per_step = cores  # one instance per host thread per step
    times = []
    pts = 0
    for i in range(args.warmup + args.steps):
        wall, pts = cpu_reference_leg(args.netlist, per_step, cores, args.tran_steps)
        if i >= args.warmup:
            times.append(wall)
    ms = 1e3 * float(np.mean(times))
    val = pts / (ms / 1e3)
    sample = f"{per_step} instances x {pts // per_step} timepoints per step, one instance per host thread"
    maps = open("/proc/self/maps").read() if os.path.exists("/proc/self/maps") else ""
    assert "libcsim_b200" not in maps, "the reference arm must not map the product library"
    print(json.dumps({
        "impl": "reference", "metric": METRIC.replace("dbmixer", args.netlist), "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_name(args):
    n = "full .TRAN" if not args.tran_steps else f"first {args.tran_steps} steps"
    return (f"{args.netlist}.sp DC operating point + transient ({n}), {args.instances_per_gpu} Monte-Carlo instances "
            f"per GPU ({args.instances_per_gpu * args.gpus} total; 65,536 at 8 GPUs), 5% uniform parameter spread, "
            f"seed {SEED}")


# ------------------------------------------------------------------------------------------
def parity_check(args, out, wave, probes, n_steps, nl):
    """Instance 0 of rank 0 (the nominal circuit) of the LAST TIMED step against the reference-pinned golden
    fixture — the fixture was generated from the unmodified reference by tests/golden/make_golden.py; nothing
    under oracle/ is executed here.  Bitwise on doubles and the pivot hash, exact on the Newton total."""
    g = np.load(os.path.join(ROOT, "tests", "golden", f"{args.netlist}_nominal.npz"))
    full = n_steps == nl.n_steps
    res = {"instance": 0, "golden": f"tests/golden/{args.netlist}_nominal.npz", "full_length": bool(full)}
    bits = lambda a: np.ascontiguousarray(a, np.float64).view(np.uint64)  # noqa: E731
    w0 = wave[0].cpu().numpy()  # [n_probes, n_out]
    stride = args.out_stride
    checks = {}
    if args.netlist == "dbmixer":
        if full and 100 % stride == 0:
            mine = w0[:, 100 // stride - 1:: 100 // stride]
            checks["wave_every_100th_of_50000"] = bool(np.array_equal(bits(mine.T), bits(g["wave_dec100"][:, probes])))
        elif stride == 1:
            k = min(n_steps, g["wave_first"].shape[0])
            checks[f"wave_first_{k}"] = bool(np.array_equal(bits(w0[:, :k].T), bits(g["wave_first"][:k][:, probes])))
    else:
        if stride == 1:
            k = min(n_steps, g["wave"].shape[0])
            checks[f"wave_first_{k}"] = bool(np.array_equal(bits(w0[:, :k].T), bits(g["wave"][:k][:, probes])))
    if full:
        want_total = int(g["iters_step"].astype(np.int64).sum())
        res["newton_total"] = int(out["iters_total"][0].item())
        res["newton_total_golden"] = want_total
        checks["newton_total"] = res["newton_total"] == want_total
        checks["pivot_hash"] = bool(out["hash"][:1].cpu().numpy().view(np.uint64)[0] == np.uint64(g["hash"]))
        xf = g["x_final"] if "x_final" in g.files else g["wave"][-1]
        checks["x_final_bits"] = bool(np.array_equal(bits(out["x"][:, 0].cpu().numpy()), bits(xf)))
    res["checks"] = checks
    res["ok"] = bool(checks) and all(checks.values())
    if not res["ok"]:
        raise SystemExit(f"bench.py: PARITY FAILURE against the golden fixture: {json.dumps(res)}")
    return res


# ------------------------------------------------------------------------------------------
def symbolic_flops(nl, plan, capi, torch):
    """Useful FP64 operations of ONE Newton iteration (structural non-zeros only) under the pivot order the
    device picks for the nominal circuit, next to the reference's dense count."""
    N = nl.n_unknowns
    rowptr, colidx = plan.pattern(capi.MODE_TRAN)
    A = np.zeros((N, N), bool)
    for i in range(N):
        A[i, colidx[rowptr[i]:rowptr[i + 1]]] = True
    # numeric pivot order of the nominal instance at its DC point (device LU)
    params = torch.from_numpy(np.ascontiguousarray(nl.params[:, None])).cuda()
    d = capi.dc(plan, params)
    dt = nl.tran["tstep"]
    setup = nl.tran_setup(n_steps=1)
    sv = torch.zeros((max(nl.n_sources, 1), 1), dtype=torch.float64, device="cuda")
    vals, rhs = capi.stamp_csr(plan, capi.MODE_TRAN, params, d["x"], d["x"], sv, dt=dt, gmin=plan.consts.tran_gmin)
    G = torch.zeros((N * N, 1), dtype=torch.float64, device="cuda")
    rows = np.repeat(np.arange(N), np.diff(rowptr))
    G[torch.from_numpy(rows * N + colidx).cuda().long(), 0] = vals[:, 0]
    _, perm, _ = capi.lu_solve(G, rhs)
    perm = perm[:, 0].cpu().numpy()
    A = A[perm]
    elim = divs = 0
    for k in range(N):
        r = [i for i in range(k + 1, N) if A[i, k]]
        c = [j for j in range(k + 1, N) if A[k, j]]
        divs += len(r)
        elim += len(r) * len(c)
        for i in r:
            A[i, c] = True
    nL = int(np.tril(A, -1).sum())
    nU = int(np.triu(A, 1).sum())
    contrib = 0
    for m in (capi.MODE_TRAN,):
        contrib = int(np.diff(plan.pattern(m)[0]).sum())
    n_mos = int((nl.elem_type == capi.ELEM_MOS).sum())
    useful = 2 * elim + divs + 2 * nL + 2 * nU + N + 40 * n_mos + 2 * contrib + 3 * N
    dense = (N * (N - 1) // 2) + 2 * sum((N - 1 - k) ** 2 for k in range(N)) + 2 * N * (N - 1) + N + 40 * n_mos
    return dict(useful=useful, dense=dense, elim_pairs=elim, divs=divs + N, nnz_L=nL, nnz_U=nU)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from circuitsimulator_b200 import capi, shard
    from circuitsimulator_b200.netlist import Netlist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    nl = Netlist(load_netlist_text(args.netlist))
    plan = nl.plan(device=local)
    B = args.instances_per_gpu
    n_steps = args.tran_steps or nl.n_steps
    N, P = nl.n_unknowns, nl.n_params
    probes = nl.plot_probes if len(nl.plot_probes) else np.arange(min(2, N), dtype=np.int32)
    setup = nl.tran_setup(probes=probes, out_stride=args.out_stride, n_steps=n_steps)
    params_host = torch.from_numpy(shard.shard_params(nl, world * B, rank, world, SEED)).pin_memory()
    params = params_host.to(dev)

    wave = torch.empty((B, len(probes), max(setup.n_out, 1)), dtype=torch.float64, device=dev)
    iters_total = torch.zeros((B,), dtype=torch.int32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    kev = []  # (before DC, between DC and transient, after transient) events of every timed step

    def one_step(timed=False):
        h, st = capi.new_diag(B, dev)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if timed else None
        if timed:
            ev[0].record()
        d = capi.dc(plan, params, h, st)
        if timed:
            ev[1].record()
        iters_total.zero_()
        t = capi.tran(plan, params, d["x"], setup, wave=wave, iters_total=iters_total, pivot_hash=h, status=st)
        if timed:
            ev[2].record()
            kev.append(ev)
        return t

    for _ in range(args.warmup):
        one_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = one_step(timed=True)
    e1.record()
    barrier()
    ms_step = e0.elapsed_time(e1) / args.steps
    dc_ms = float(np.mean([ev[0].elapsed_time(ev[1]) for ev in kev]))      # launch duration of the DC kernel
    tran_ms = float(np.mean([ev[1].elapsed_time(ev[2]) for ev in kev]))    # ... of the fused transient kernel
    clocks = sampler.stop() if rank == 0 else None
    tms = torch.tensor([ms_step], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_step = float(tms.item())
    value = world * B * n_steps / (ms_step / 1e3)
    newton_total = int(out["iters_total"].sum().item())
    status_or = int(out["status"].max().item())

    # ---- parity of the run that was just timed: instance 0 of rank 0's shard is the nominal circuit, whose
    # reference-pinned run is the golden fixture (tests/golden/<netlist>_nominal.npz, made by make_golden.py from
    # the unmodified reference): final state and pivot hash bit for bit, Newton total, every 100th waveform row
    parity = None
    if rank == 0:
        parity = parity_check(args, out, wave, probes, n_steps, nl)

    # ---- strong scaling: a FIXED total (65,536 instances) split over the N GPUs, one timed run, decimated output
    strong = None
    if args.strong_total and args.strong_total % world == 0:
        Bs_ = args.strong_total // world
        if Bs_ == B:
            strong = {"total_instances": args.strong_total, "instances_per_gpu": Bs_, "value": value, "unit": UNIT,
                      "ms": ms_step, "out_stride": args.out_stride, "note": "same shard size as the main run"}
        else:
            ssetup = nl.tran_setup(probes=probes, out_stride=args.strong_out_stride, n_steps=n_steps)
            sparams = torch.from_numpy(shard.shard_params(nl, args.strong_total, rank, world, SEED)).to(dev)
            swave = torch.empty((Bs_, len(probes), max(ssetup.n_out, 1)), dtype=torch.float64, device=dev)
            stot = torch.zeros((Bs_,), dtype=torch.int32, device=dev)

            def strong_step():
                h, st = capi.new_diag(Bs_, dev)
                d = capi.dc(plan, sparams, h, st)
                stot.zero_()
                return capi.tran(plan, sparams, d["x"], ssetup, wave=swave, iters_total=stot, pivot_hash=h, status=st)

            strong_step()  # warm-up (kernel variants, clocks)
            barrier()
            q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            q0.record()
            sout = strong_step()
            q1.record()
            barrier()
            sms = torch.tensor([q0.elapsed_time(q1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(sms, op=dist.ReduceOp.MAX)
            strong = {"total_instances": args.strong_total, "instances_per_gpu": Bs_,
                      "value": args.strong_total * n_steps / (float(sms.item()) / 1e3), "unit": UNIT,
                      "ms": float(sms.item()), "out_stride": args.strong_out_stride,
                      "status_or": int(sout["status"].max().item()),
                      "note": "one timed run after one warm-up, device-resident, max over ranks"}
            del sparams, swave, stot, sout

    # ---- e2e: the host-buffer C-ABI call (csim_batch_tran): H2D params + tables, D2H waveforms + diagnostics
    e2e = None
    if not args.no_e2e:
        batch = capi.Batch(plan, B)

        def pinned(shape, dtype):
            return torch.empty(shape, dtype=dtype).pin_memory().numpy()

        hout = dict(x_dc=pinned((N, B), torch.float64), x_final=pinned((N, B), torch.float64),
                    wave=pinned((B, len(probes), max(setup.n_out, 1)), torch.float64),
                    iters_total=pinned((B,), torch.int32), iters_step=None,
                    hash=pinned((B,), torch.int64).view(np.uint64), status=pinned((B,), torch.int32).view(np.uint32))
        ph = params_host.numpy()

        def e2e_step():
            batch.set_params(ph)
            batch.tran(setup, out=hout)

        e2e_step()  # warm-up (allocations, first-touch)
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 2))
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        wall = (time.perf_counter() - t0) / n_e2e
        tw = torch.tensor([wall], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        h2d = ph.nbytes + setup.table.nbytes
        d2h = sum(v.nbytes for v in hout.values() if v is not None)
        e2e = {"value": world * B * n_steps / float(tw.item()), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "api": "csim_batch_set_params + csim_batch_tran (host buffers)"}
        # the host-API result must equal the device-API result
        assert np.array_equal(hout["iters_total"], out["iters_total"].cpu().numpy()), "e2e result differs"
        batch.close()

    # ---- final gather of the probe waveforms (outside the metric, reported separately)
    gather_ms = None
    if world > 1:
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dec = wave[:, :, :: max(1, wave.shape[2] // 5000)].contiguous()  # decimated to <= 5000 points for the gather
        shard.gather_instances(dec[:, :, :8].contiguous(), dst=0, dim=0)  # untimed: NCCL sets its P2P channels up lazily
        torch.cuda.synchronize()
        g0.record()
        allw = shard.gather_instances(dec, dst=0, dim=0)
        alli = shard.gather_instances(out["iters_total"], dst=0)
        g1.record()
        torch.cuda.synchronize()
        gather_ms = g0.elapsed_time(g1)
        if rank == 0:
            assert allw.shape[0] == world * B and alli.shape[0] == world * B

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline leg: the standalone stamp kernel (HBM-bound), inputs larger than L2
    peaks, peak_kind = measured_peaks()
    Bs = args.stamp_instances
    nnz = plan.pattern(capi.MODE_TRAN)[1].size
    reps = (Bs + B - 1) // B
    sp = params.repeat(1, reps)[:, :Bs].contiguous()
    sx = (torch.rand((N, Bs), dtype=torch.float64, device=dev) * 3.0)
    sxp = sx + 0.01
    ssv = torch.ones((max(nl.n_sources, 1), Bs), dtype=torch.float64, device=dev)
    svals = torch.empty((nnz, Bs), dtype=torch.float64, device=dev)
    srhs = torch.empty((N, Bs), dtype=torch.float64, device=dev)
    for _ in range(3):
        capi.stamp_csr(plan, capi.MODE_TRAN, sp, sx, sxp, ssv, dt=setup.dt, gmin=1e-6, out=(svals, srhs))
    torch.cuda.synchronize()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    nrep = 10
    s0.record()
    for _ in range(nrep):
        capi.stamp_csr(plan, capi.MODE_TRAN, sp, sx, sxp, ssv, dt=setup.dt, gmin=1e-6, out=(svals, srhs))
    s1.record()
    torch.cuda.synchronize()
    stamp_ms = s0.elapsed_time(s1) / nrep
    # Algorithmic bytes per instance = the DISTINCT doubles one TRAN assembly has to read (device parameters,
    # the MOSFET terminals of x, the companion-model entries of x_prev, one value per source) and write
    # (CSR values + RHS) — counted by the plan (csim_plan_stamp_info).  SURVEY.md §8(d)'s upper figure
    # 8*(P+N+N+nnz+N) also counts source-waveform parameters and x/x_prev entries no stamp reads; it is
    # reported next to it, the roofline fraction uses the smaller, exact one.
    sinfo = plan.stamp_info(capi.MODE_TRAN, True)
    bytes_per_inst = sinfo["bytes_read"] + sinfo["bytes_written"]
    achieved = bytes_per_inst * Bs / (stamp_ms / 1e3) / 1e9
    # dram bytes read + written per launch, from the committed ncu --set full capture of this kernel at this
    # launch size (profiles/stamp_traffic.json, written by tools/stamp_traffic.py); null when it does not apply
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "stamp_traffic.json")))
        if sinfo["active"] and tj["netlist"] == args.netlist and int(tj["instances"]) == Bs:
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
    except Exception:
        pass
    roofline_stamp = {"kernel": "csim_stamp_spec<TRAN>" if sinfo["active"] else "stamp_csr_kernel<TRAN>", "bound": "hbm",
                      "achieved": achieved, "peak": peaks["hbm_gbs"],
                      "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_src,
                      "algorithmic_bytes_per_launch": bytes_per_inst * Bs, "peak_kind": peak_kind,
                      "bytes_per_instance": bytes_per_inst, "bytes_per_instance_survey_formula": 8 * (P + N + N + nnz + N),
                      "instances": Bs, "ms": stamp_ms, "in_timed_region": False,
                      "l2": "inputs+outputs %.2f GB > 126 MB L2" % (bytes_per_inst * Bs / 1e9)}
    del sp, sx, sxp, ssv, svals, srhs

    # ---- roofline of the DOMINANT kernel of the timed region: the fused transient (FP64 pipe, not HBM: state stays
    # on chip).  achieved = useful FP64 operations of one launch (structural non-zeros of one Newton iteration x the
    # Newton iterations the launch ran) / mean launch duration (CUDA events around the launch on the
    # launching stream, inside the timed loop); peak = separate DMUL+DADD throughput measured in THIS run
    # (csim_dev_fp64_peak; FMA is not allowed on the reference's expressions).
    fl = symbolic_flops(nl, plan, capi, torch)
    fp64_peak = capi.fp64_peak(local)
    fused_kind = plan.spec_info().get("kind", 0)
    kname = {3: "csim_tran_tpi (NVRTC, thread per instance)", 2: "csim_tran_spec (NVRTC, warp per instance)"}.get(
        fused_kind, "tran_kernel (generic)")
    if fused_kind == 3 and "warp-pair variant" in plan.spec_info().get("log", ""):
        kname = "csim_tran_tpi, warp-pair variant (NVRTC, two warps per 32 instances)"
    tran_tflops = newton_total * fl["useful"] / (tran_ms / 1e3) / 1e12
    tj = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "tran_traffic.json")))
        if (tj.get("netlist") != args.netlist or int(tj.get("instances", -1)) != B or int(tj.get("timepoints", -1)) != n_steps
                or int(tj.get("out_stride", -1)) != args.out_stride):
            tj = None  # the capture is of another launch size
    except Exception:
        tj = None
    roofline = {"kernel": kname, "bound": "fp64", "achieved": tran_tflops, "peak": fp64_peak,
                "unit": "TFLOP/s", "frac": tran_tflops / fp64_peak if fp64_peak else None,
                "traffic": tj["dram_bytes_per_launch"] if tj else None, "traffic_source": tj["source"] if tj else None,
                "peak_kind": "separate DMUL+DADD (no FMA), measured in this run by csim_dev_fp64_peak",
                "ms_per_launch": tran_ms, "share_of_step": tran_ms / ms_step, "dc_kernel_ms": dc_ms,
                "useful_flops_per_newton_iter": fl["useful"], "newton_iters_per_launch": newton_total,
                "flops_per_newton_iter": fl, "newton_iters_per_timepoint": newton_total / (B * n_steps),
                "hbm_bytes_per_timepoint": 8 * len(probes) / args.out_stride,
                "dense_equiv_tflops": newton_total * fl["dense"] / (tran_ms / 1e3) / 1e12,
                "note": "dense_equiv counts what the reference executes (zeros included); it is a work measure, not a utilisation"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            from oracle import oraclelib as O
            if O.ref_available():
                cores = O.ref_hardware_threads()
                n_inst = args.cpu_sample or 2 * cores
                wall, pts = cpu_reference_leg(args.netlist, n_inst, cores, args.tran_steps)
                cpu = {"value": pts / wall, "unit": UNIT, "cores": cores, "kind": "reference",
                       "sample": f"{n_inst} instances of the same Monte-Carlo batch x {pts // n_inst} timepoints, "
                                 f"unmodified reference (oracle/_ref, g++ -O3 -ffp-contract=off), one instance per "
                                 f"host thread, {wall:.1f} s wall"}
        except Exception as ex:  # the baseline is a reported number, never a reason to lose the bench line
            cpu = {"value": None, "unit": UNIT, "cores": None, "kind": "reference", "sample": f"failed: {ex}"}

    line = {
        "metric": METRIC.replace("dbmixer", args.netlist), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args), "netlist": args.netlist, "instances_per_gpu": B,
                   "timepoints": n_steps, "probes": [int(p) for p in probes], "out_stride": args.out_stride,
                   "l2": "no flush needed: state is on-chip, per-step outputs stream to a %.2f GB buffer > L2"
                         % (wave.numel() * 8 / 1e9),
                   "parallelism": f"instances sharded over {world} GPU(s), no data-path collective"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": 2 * args.steps,
        "roofline": roofline, "roofline_stamp": roofline_stamp, "cpu_baseline": cpu, "parity_check": parity,
        "strong": strong,
        "newton_iterations": newton_total, "status_or": status_or, "gather_ms": gather_ms,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    import __graft_entry__ as g
    if args.impl == "reference":
        g.build(load_library=False)  # the reference arm never maps libcsim_b200.so
        run_reference(args)
    else:
        g.build()
        run_ours(args)


if __name__ == "__main__":
    main()
