#!/usr/bin/env python
"""bench.py -- BN254 Groth16 prove @ 2^22 constraints on N B200s (BASELINE.json configs[3]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--log-n 22] [--no-extras] [--no-cpu]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the proving hot path (computeH + the five MSMs + window Horner on the host) over one solved
instance of the MiMC-shaped chain R1CS (real witness, trapdoor setup; tests/helpers.py ChainCase).  `value` is timed with
the inputs resident in HBM (g16_prove_bn254_dev); `e2e` goes through the plugin call with pinned HOST buffers
(g16_prove_bn254: H2D of W, a, b, c inside the timed region, the five result points read back); `e2e_pageable` is the
same call on plain pageable arrays (a Go caller's slices); `e2e_wires` ships only the wire vector.  N > 1: ONE process
drives all N GPUs through one multi-GPU g16 context -- the reference surface is a single in-process call -- so under
torchrun rank 0 makes the calls and the other ranks only keep the barriers and the max-over-ranks reduction.
Before anything is timed, all five points of the instance are checked against the oracle (`parity` in the line).
oracle/ is used only as that checker, for the instance's inputs, and in the cpu_baseline / --impl reference legs.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

FR_TOP_MASK = 0x0FFFFFFFFFFFFFFF   # any 252-bit value is a valid (Montgomery) Fr / scalar (both curves: r > 2^252)
IMAD_EQ_PER_FP_MUL = 272           # (2*8^2 + 8) limb-MACs x 2; 8-limb fields: BN254 Fp/Fr, BLS12-377 Fr
IMAD_EQ_PER_FP_MUL_12 = 600        # same formula for the 12-limb BLS12-377 Fp
MADD_MULS = {"g1": 10, "g2": 30}
CURVE_NAMES = {"bn254": "bn254", "bls12_377": "bls12_377"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.stop, self.idx = [], threading.Event(), gpu_index
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.idx)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for nme, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------------ workload
def chain_shape(nc, power=5):
    """infinity masks with the MiMC-chain shape (pow5 rounds, the u = t*t wires never appear in R; pow17
    rounds (BLS12-377, encrypt.go:137-143) are 5 constraints of which the t^16 wire only appears in L)."""
    nw = nc + 3
    inf_a = np.zeros(nw, dtype=np.uint8)
    inf_b = np.zeros(nw, dtype=np.uint8)
    inf_a[1] = 1
    if power == 5:
        inf_b[4::3] = 1
    else:
        inf_b[7::5] = 1
    inf_b[1] = 1
    return nw, inf_a, inf_b


def rand_fr_cuda(torch, n, gen, dev):
    x = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n, 4), dtype=torch.int64, device=dev, generator=gen)
    x[:, 3] &= FR_TOP_MASK
    return x


def build_inputs(torch, ctx, nc, dev, seed=20260422):
    """device-side synthetic key + solution; returns host key arrays and device/pinned solution buffers."""
    n = 1
    while n < nc:
        n <<= 1
    nw, inf_a, inf_b = chain_shape(nc, 5 if ctx.curve == "bn254" else 17)
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed)
    na, nb, nk, nz = int((inf_a == 0).sum()), int((inf_b == 0).sum()), nw - 2, n - 1

    def g1_bases(cnt):
        k = rand_fr_cuda(torch, cnt, gen, dev)
        out = torch.empty((cnt, ctx.g1w), dtype=torch.int64, device=dev)
        ctx.g1_fixed_base_mul_dev(k.data_ptr(), cnt, out.data_ptr())
        return out.cpu().numpy().view(np.uint64)

    key = dict(domain_n=n, n_wires=nw, inf_a=inf_a, inf_b=inf_b,
               g1_a=g1_bases(na), g1_b=g1_bases(nb), g1_k=g1_bases(nk), g1_z=g1_bases(nz))
    k2 = rand_fr_cuda(torch, nb, gen, dev)
    out2 = torch.empty((nb, ctx.g2w), dtype=torch.int64, device=dev)
    ctx.g2_fixed_base_mul_dev(k2.data_ptr(), nb, out2.data_ptr())
    key["g2_b"] = out2.cpu().numpy().view(np.uint64)
    del k2, out2
    # solution: W random, a, b random, c = a*b (satisfied system => the quotient exists)
    W = rand_fr_cuda(torch, nw, gen, dev)
    a = rand_fr_cuda(torch, n, gen, dev)
    b = rand_fr_cuda(torch, n, gen, dev)
    a[nc:] = 0
    b[nc:] = 0
    a_h, b_h = a.cpu().numpy().view(np.uint64), b.cpu().numpy().view(np.uint64)
    c_h = ctx.fr_mul_vec(a_h, b_h)
    c = torch.from_numpy(c_h.view(np.int64)).to(dev)
    sol_dev = dict(W=W, a=a, b=b, c=c)
    sol_host = {k: v.cpu().pin_memory() for k, v in (("W", W), ("a", a[:nc]), ("b", b[:nc]), ("c", c[:nc]))}
    return key, sol_dev, sol_host, n, nw


# ------------------------------------------------------------------------------------------------ the BN254 instance
BENCH_SEED = 22
POINT_NAMES = ["ar", "bs1", "krs_k", "krs_hz", "bs2"]


def oracle_for(curve):
    """the C oracle of the curve (the checker, and the timed CPU prover)"""
    if curve == "bls12_377":
        from oracle.bls12_377 import cpu as oc
    else:
        from oracle import cpu as oc
    return oc


def chain_case(nc, g1_mul=None, g2_mul=None, curve="bn254"):
    """BASELINE config 4 (BN254, pow5 rounds) / config 5 (BLS12-377, pow17 rounds): the MiMC-shaped chain at `nc`
    constraints with a trapdoor setup (tests/helpers.py ChainCase: witness, a, b, c and the setup scalars A(t), B(t), K, Z
    from oracle/c; bases = [scalar]G from g1_mul / g2_mul -- the library's fixed-base kernel in the GPU arm, oracle/c's
    in the reference arm: same seed, identical key)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import ChainCase
    return ChainCase(nc, seed=BENCH_SEED, g1_mul=g1_mul, g2_mul=g2_mul, oc=oracle_for(curve))


def workload_config(case, log_n, curve="bn254"):
    pk = case.pk
    if curve == "bls12_377":
        return {"workload": f"groth16_bls12_377_prove_2p{log_n}_mimc17_chain", "constraints": case.nc, "domain": case.n, "wires": case.nw,
                "msm_sizes": {"g1_a": int(len(pk.g1_a)), "g1_b": int(len(pk.g1_b)), "g1_k": int(len(pk.g1_k)),
                              "g1_z": int(len(pk.g1_z)), "g2_b": int(len(pk.g2_b))},
                "instance": f"MiMC-shaped pow17 chain (5 constraints per round, std/hash/mimc/encrypt.go:137-143), trapdoor setup, xoshiro seed {BENCH_SEED}",
                "l2": "inputs larger than L2"}
    return {"workload": f"groth16_bn254_prove_2p{log_n}_mimc_chain", "constraints": case.nc, "domain": case.n,
            "wires": case.nw,
            "msm_sizes": {"g1_a": int(len(pk.g1_a)), "g1_b": int(len(pk.g1_b)), "g1_k": int(len(pk.g1_k)),
                          "g1_z": int(len(pk.g1_z)), "g2_b": int(len(pk.g2_b))},
            "instance": f"MiMC-shaped pow5 chain (3 constraints per round), trapdoor setup, xoshiro seed {BENCH_SEED}",
            "l2": "inputs larger than L2"}


def cpu_prove(case, ocpu):
    """the reference's CPU prover schedule (backend/groth16/bn254/prove.go:141-150,306-320) as restated in oracle/c:
    computeH on all cores, then the five MultiExps; returns (ms, the five points, h)."""
    pk = case.pk
    wa, wb, wk = case.W[~case.inf_a], case.W[~case.inf_b], case.W[2:]
    t0 = time.perf_counter()
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    pts = np.concatenate([ocpu.msm_g1(pk.g1_a, wa), ocpu.msm_g1(pk.g1_b, wb), ocpu.msm_g1(pk.g1_k, wk),
                          ocpu.msm_g1(pk.g1_z, h[: case.n - 1]), ocpu.msm_g2(pk.g2_b, wb)])
    return (time.perf_counter() - t0) * 1e3, pts, h


def parity_report(got, want, method, curve="bn254"):
    got, want = np.asarray(got, dtype=np.uint64).reshape(-1), np.asarray(want, dtype=np.uint64).reshape(-1)
    w = 12 if curve == "bls12_377" else 8      # uint64 limbs of a G1 affine point; G2 = 2 w
    bounds = [0, w, 2 * w, 3 * w, 4 * w, 6 * w]
    bad = [POINT_NAMES[i] for i in range(5) if not np.array_equal(got[bounds[i]:bounds[i + 1]], want[bounds[i]:bounds[i + 1]])]
    return {"checked": POINT_NAMES, "bytes": int(got.size * 8), "mismatches": len(bad), "mismatched": bad, "method": method}


def run_reference(args):
    """--impl reference: the CPU prover (oracle/c restatement of backend/groth16/bn254/prove.go; gnark-crypto itself
    needs Go, which neither machine has) on the SAME instance and config as the GPU arm, all host cores, every timed
    step a full 2^log_n prove.  Touches no CUDA and does not load libg16b200.so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ocpu = oracle_for(args.curve)
    t_build = time.perf_counter()
    case = chain_case(1 << args.log_n, curve=args.curve)   # key from oracle/c's batch fixed-base multiplication
    t_build = time.perf_counter() - t_build
    want = None
    times, budget = [], args.cpu_budget_s
    warm = min(args.warmup, 1)                  # one untimed prove warms the page cache / thread pool; a CPU needs no more
    for it in range(warm + args.steps):
        ms, pts, h = cpu_prove(case, ocpu)
        if want is None:
            want = case.expected_points(h)      # exponent side: [sum w_i s_i]G with the setup's trapdoor scalars
        par = parity_report(pts, want, "trapdoor identity on the oracle prover's five points", curve=args.curve)
        if par["mismatches"]:
            raise SystemExit(f"bench.py --impl reference: CPU prover disagrees with the trapdoor identity: {par}")
        if it >= warm:
            times.append(ms)
            if sum(times) / 1e3 > budget:       # bounded: stop once the timed region exceeds the budget
                break
    v = sum(times) / len(times)
    cores = ocpu.threads()
    sample = (f"oracle/c restatement of the gnark-crypto CPU prover (not gnark-crypto itself: no Go toolchain, no ADX/BMI2 "
              f"assembly), {len(times)} full prove(s) at 2^{args.log_n} constraints on {cores} threads; key built by oracle/c in {t_build:.0f} s")
    line = {"impl": "reference", "metric": f"groth16_{args.curve}_prove_ms", "value": v, "unit": "ms", "n_gpus": args.gpus,
            "steps": len(times), "steps_requested": args.steps, "steps_run": len(times), "warmup": warm, "warmup_requested": args.warmup,
            "ms_per_step": v, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32x8-montgomery" if args.curve == "bn254" else "u32x12-montgomery",
            "data": "synthetic", "config": workload_config(case, args.log_n, args.curve), "parity": par,
            "cpu_baseline": {"value": v, "unit": "ms", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ GPU arm
class Ranks:
    """torchrun plumbing.  The product drives all N GPUs from ONE process through one multi-GPU g16 context (the
    reference surface is a single in-process call); under torchrun rank 0 does that while the other ranks keep the
    contract's barriers and the max-over-ranks reduction company."""

    def __init__(self, torch):
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            self.dist = dist
            dev = torch.device(f"cuda:{self.local}")
            torch.cuda.set_device(dev)
            # NCCL prints its version banner on stdout when the communicator comes up; rank 0 must print ONE JSON line
            # there, so stdout points at stderr until then
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=dev)
                warm = torch.zeros(1, device=dev)
                dist.all_reduce(warm)
                torch.cuda.synchronize(dev)
                # host-side rendezvous: a rank that waits inside an NCCL collective keeps a spinning kernel on its GPU,
                # which time-slices with the kernels rank 0's context runs there (measured: 2x slower steps)
                self.host_group = dist.new_group(backend="gloo")
            finally:
                sys.stdout.flush()
                os.dup2(saved, 1)
                os.close(saved)

    def barrier(self):
        if self.dist:
            self.dist.barrier(group=self.host_group)   # wait on the CPU until every rank is here ...
            self.dist.barrier()                        # ... then the NCCL barrier returns at once

    def max_ms(self, ms):
        if not self.dist:
            return ms
        self.dist.barrier(group=self.host_group)
        t = self.torch.tensor([ms], device=f"cuda:{self.local}")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


def timed(torch, ranks, ext, fn, steps, warmup):
    """W warm-up calls, then K calls bracketed by barrier + synchronize and CUDA events on the library's stream (the
    calls are synchronous: the closing event is recorded after the last one has returned); max over ranks."""
    for _ in range(warmup):
        fn()
    ranks.barrier()
    torch.cuda.synchronize()
    ms, last = 0.0, None
    if fn is not None and ext is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for _ in range(steps):
            last = fn()
        e1.record(ext)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    ranks.barrier()
    return ranks.max_ms(ms) / steps, last


def idle_rank(torch, ranks, legs, steps):
    for _ in range(legs):
        ranks.barrier()
        torch.cuda.synchronize()
        ranks.barrier()
        ranks.max_ms(0.0)
    ranks.close()


def run_ours(args):
    import torch
    from gnark_icicle_b200 import lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the accelerated path has no CPU fallback")
    if args.curve != "bn254":
        return run_ours_synthetic(args)
    ranks = Ranks(torch)
    world = ranks.world if ranks.world > 1 else max(args.gpus, 1)   # without torchrun: the same single process, N devices
    LEGS = 4                                     # value, e2e (pinned), e2e (pageable), e2e (wires only)
    if ranks.rank != 0:
        return idle_rank(torch, ranks, LEGS, args.steps)
    from oracle import cpu as ocpu               # the checker: expected points, and the cpu_baseline leg
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    devices = list(range(world))
    ctx = lib.Context(devices if world > 1 else 0)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    ctx.set("msm_ba_rounds", args.ba_rounds)
    if args.g2_weight:
        ctx.set("g2_weight", args.g2_weight)
    nc = 1 << args.log_n
    case = chain_case(nc, g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul)
    n, nw, pk = case.n, case.nw, case.pk
    cfg = workload_config(case, args.log_n)
    t_up = time.perf_counter()
    handle = ctx.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, nw)
    t_up = time.perf_counter() - t_up
    placement = [ctx.pk_placement(handle, r) for r in range(world)]

    # ---- parity first: every one of the five points of THIS instance against the exponent-side
    # expectation [sum_i w_i s_i]G, s = the setup's trapdoor scalars, h from the CPU oracle's computeH
    h_cpu = ocpu.compute_h(case.a, case.b, case.c, n)
    want = case.expected_points(h_cpu)
    pinned = {k: torch.from_numpy(v.view(np.int64)).pin_memory() for k, v in (("W", case.W), ("a", case.a), ("b", case.b), ("c", case.c))}
    hW, ha, hb, hc = (pinned[k].numpy().view(np.uint64) for k in ("W", "a", "b", "c"))
    pts, _ = ctx.prove(handle, hW, ha, hb, hc)
    parity = parity_report(pts, want, "trapdoor identity [sum w_i s_i]G on all five points of the benchmarked instance, h from the "
                                      "CPU oracle's computeH; checked before timing" + (f", on the points combined over {world} GPUs" if world > 1 else ""))
    if parity["mismatches"]:
        raise SystemExit(f"bench.py: GPU prove disagrees with the oracle: {parity}")

    # ---- device-resident inputs (value)
    Wd = pinned["W"].to(dev)
    abc_master = torch.zeros((3, n, 4), dtype=torch.int64, device=dev)
    for i, k in enumerate(("a", "b", "c")):
        abc_master[i, :nc] = pinned[k].to(dev)
    abc_work = torch.empty_like(abc_master)

    def step_dev():
        with torch.cuda.stream(ext):                    # a, b, c are clobbered by the transform pipeline
            abc_work.copy_(abc_master, non_blocking=True)
        return ctx.prove_dev(handle, Wd.data_ptr(), abc_work[0].data_ptr(), abc_work[1].data_ptr(), abc_work[2].data_ptr())

    def step_e2e():
        return ctx.prove(handle, hW, ha, hb, hc)        # pinned host buffers in, 5 points out

    pW, pa, pb, pc = case.W, case.a, case.b, case.c     # plain numpy = pageable memory: what a Go caller passes

    def step_pageable():
        return ctx.prove(handle, pW, pa, pb, pc)

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import chain_csr
    from oracle.prng import SEED_BASE
    cs = chain_csr(nc, SEED_BASE + 2000 + BENCH_SEED)
    r1 = ctx.r1cs_upload(cs.nb_constraints, cs.nb_wires, cs.coefficients, (cs.L, cs.R, cs.O))

    def step_wires():
        return ctx.prove_wires(handle, r1, hW)          # only the wire vector crosses PCIe

    l0 = lib.kernel_launch_count(args.curve)
    with ClockSampler(0) as clk:
        ms_dev, (pts_dev, st) = timed(torch, ranks, ext, step_dev, args.steps, args.warmup)
    launches = (lib.kernel_launch_count(args.curve) - l0) * args.steps // (args.steps + args.warmup)
    ms_e2e, (pts_e2e, st_e) = timed(torch, ranks, ext, step_e2e, args.steps, args.warmup)
    ms_pg, (pts_pg, st_pg) = timed(torch, ranks, ext, step_pageable, args.steps, args.warmup)
    ms_w, (pts_w, st_w) = timed(torch, ranks, ext, step_wires, args.steps, args.warmup)
    for name, p in (("device-resident", pts_dev), ("pinned e2e", pts_e2e), ("pageable e2e", pts_pg), ("wires-only e2e", pts_w)):
        if not np.array_equal(p, want):
            raise SystemExit(f"bench.py: {name} prove differs from the oracle's expectation")
    parity["legs_checked"] = ["value", "e2e", "e2e_pageable", "e2e_wires"]
    ctx.r1cs_free(r1)
    ranks.close()

    # per-kernel durations for the roofline: one extra step with the MSMs serialised (with overlap on, kernels of
    # different MSMs share the SMs and their individual event times are not meaningful)
    ctx.set("streams", 1)
    step_dev()
    _, st = step_dev()
    ctx.set("streams", 5)
    imad, wide = ctx.imad_peak()
    wide_x, fr_mul, fp_mul = ctx.pipe_rates()
    fp_eq = IMAD_EQ_PER_FP_MUL
    # dominant kernel: k_accumulate<G1> (4 launches per step and GPU); algorithmic IMAD-equivalents per launch =
    # entries (points x windows) x 10 Fp multiplications per XYZZ mixed add x 272.  On a group the
    # per-MSM time is the slowest rank's and the work is the whole job's: the fraction is of the N-GPU aggregate peak.
    per_rank_entries = []
    for r in range(world):
        per_rank_entries.append([(hi - lo) for lo, hi in placement[r]])
    g1_ops = sum(st["msm_n"][i] * st["msm_windows"][i] for i in range(4)) * MADD_MULS["g1"] * fp_eq
    g1_ms = sum(st["ms_accumulate"][:4])
    g2_ops = st["msm_n"][4] * st["msm_windows"][4] * MADD_MULS["g2"] * fp_eq
    g2_ms = st["ms_accumulate"][4]
    peak = imad * world
    ach = g1_ops / (g1_ms * 1e-3) / 1e12
    # what the stage EXECUTES since the batch-affine rounds (csrc/msm_ba.cuh): an affine pair addition is 6 multiplications
    # (3 for the shared inversion), R rounds leave 1/2^R of the entries to the 10-multiplication XYZZ addition
    ba_r = args.ba_rounds
    mults_exec = (1 - 0.5 ** ba_r) * 6 + 0.5 ** ba_r * 10 if ba_r > 0 else 10.0
    roof = {"bound": "int32-imad", "kernel": "bucket accumulation of the four G1 MSMs: k_ba_fwd / k_ba_bwd rounds + k_accumulate<G1>",
            "achieved": ach, "peak": peak / 1e12,
            "unit": "TIMAD/s", "frac": ach / (peak / 1e12), "traffic": None,
            "frac_definition": "SURVEY 8(d): entries x 10 Fp mul x 272 IMAD-eq (the reference algorithm's work) / stage time / measured IMAD peak",
            "mults_per_entry_executed": mults_exec, "frac_executed": ach / (peak / 1e12) * mults_exec / 10.0,
            "batch_affine_rounds": ba_r,
            "peak_source": "g16_imad_peak microbenchmark, same run (independent mad.lo.u32 chains on all SMs)" + (f" x {world} GPUs" if world > 1 else ""),
            "imad_wide_peak_tops": wide / 1e12, "imad_wide_carry_peak_tops": wide_x / 1e12,
            "field_mul_peak_g_per_s": fp_mul / 1e9, "field_mul_achieved_g_per_s": g1_ops / fp_eq / (g1_ms * 1e-3) / 1e9 / world,
            "frac_of_multiplier_ceiling": g1_ops / fp_eq / (g1_ms * 1e-3) / fp_mul / world,
            "g2": {"kernel": "bucket accumulation of the G2 MSM: k_ba_fwd / k_ba_bwd rounds + k_accumulate<G2>",
                   "achieved": g2_ops / (g2_ms * 1e-3) / 1e12, "frac": g2_ops / (g2_ms * 1e-3) / peak,
                   "frac_definition": "entries x 30 Fp mul x 272 IMAD-eq / stage time / measured IMAD peak (SURVEY 8(d))",
                   "mults_per_entry_executed": ((1 - 0.5 ** ba_r) * 15 + 0.5 ** ba_r * 28) if ba_r > 0 else 28.0,
                   "frac_executed": g2_ops / (g2_ms * 1e-3) / peak * (((1 - 0.5 ** ba_r) * 15 + 0.5 ** ba_r * 28) if ba_r > 0 else 28.0) / 30.0},
            "share_of_step": (g1_ms + g2_ms) / st["ms_total"],
            "note": "stage times from one extra step with the five MSMs serialised (streams=1); the timed steps overlap them, which is "
                    "where the batch-affine rounds pay: their gathers and inversion trees hide under other MSMs' multiplications "
                    "(serialised sum 67 ms, overlapped step 56 ms)"
                    + ("; on a group: slowest rank's time per MSM against the whole job's work" if world > 1 else "")}
    if args.log_n == 22 and world == 1 and ba_r == 4:
        roof["traffic"] = 20.8e9
        roof["traffic_source"] = ("profiles/r02_ba_accumulate_2p22_ncu_raw.csv (ncu --set full): dram__bytes_read + write summed over the nine kernels "
                                  "of one 2^22-point G1 accumulation stage (4 k_ba_fwd + 4 k_ba_bwd + k_accumulate): 19.1 GB read + 5.2 GB written "
                                  "before the 64-byte fill hint in k_ba_fwd (-3.5 GB), against 3.71 GB for the direct XYZZ kernel of round 1 "
                                  "(54.5 M entries x 68 B): the rounds trade multiplications for memory traffic")
    pk_, src = peaks()
    # HBM view of the same kernel: 64 B affine point + 4 B sorted index per bucket entry; `traffic` is
    # dram__bytes_read+write per launch of the 2^22-point launch from the ncu --set full capture in profiles/
    g1_bytes = sum(st["msm_n"][i] * st["msm_windows"][i] for i in range(4)) * (8 * ctx.g1w + 4)
    roof["hbm"] = {"bound": "hbm", "achieved": g1_bytes / (g1_ms * 1e-3) / 1e9, "peak": pk_["hbm_gbs"] * world, "unit": "GB/s",
                   "frac": g1_bytes / (g1_ms * 1e-3) / 1e9 / (pk_["hbm_gbs"] * world), "peak_source": src,
                   "algorithmic_bytes_per_launch_2p22": (1 << 22) * 13 * 68,
                   "note": "the accumulate kernel is bound by the integer multiplier, not HBM: it moves <10 % of the copy bandwidth"}
    ntt_bytes = 7 * 64 * n
    if st["ms_compute_h"] > 0:
        roof["compute_h"] = {"bound": "hbm", "achieved": ntt_bytes / (st["ms_compute_h"] * 1e-3) / 1e9, "peak": pk_["hbm_gbs"] * world,
                             "unit": "GB/s", "frac": ntt_bytes / (st["ms_compute_h"] * 1e-3) / 1e9 / (pk_["hbm_gbs"] * world),
                             "peak_source": src, "ms": st["ms_compute_h"],
                             "note": "7 transforms x 64n B algorithmic; the butterflies are IMAD-bound at 254 bits"
                                     + ("; distributed: slowest rank, exchanges and their waits included" if world > 1 else "")}
    par = "1 GPU" if world == 1 else (f"{world} GPUs behind one g16 context: key placed on a cost line (G2 point = 3 G1 points), computeH split at one "
                                      "stage with 3 peer-store exchanges over NVLink, partial points summed on the host")
    line = {"metric": "groth16_bn254_prove_ms", "value": ms_dev, "unit": "ms", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32x8-montgomery", "data": "synthetic", "config": cfg, "parallelism": par, "parity": parity,
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": int(st_e["h2d_bytes"]),
                    "d2h_bytes_per_step": int(st_e["d2h_bytes"]), "host_memory": "pinned"},
            "e2e_pageable": {"value": ms_pg, "unit": "ms", "h2d_bytes_per_step": int(st_pg["h2d_bytes"]), "d2h_bytes_per_step": int(st_pg["d2h_bytes"]),
                             "note": "same call with plain (pageable) numpy arrays: what a Go caller's heap slices are"},
            "e2e_wires": {"value": ms_w, "unit": "ms", "h2d_bytes_per_step": int(st_w["h2d_bytes"]), "d2h_bytes_per_step": int(st_w["d2h_bytes"]),
                          "note": "g16_prove_bn254_wires: a, b, c evaluated on the device from resident constraint matrices (8(f) N1)"},
            "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roof, "key_upload_s": t_up,
            "placement": {f"gpu{r}": {POINT_NAMES[i]: list(placement[r][i]) for i in range(5) if placement[r][i][1] > placement[r][i][0]} for r in range(world)},
            "stages_ms": {"serialised_total": st["ms_total"], "compute_h": st["ms_compute_h"], "msm": st["ms_msm"], "accumulate": st["ms_accumulate"],
                          "host_epilogue": st["ms_host"], "msm_c": st["msm_c"], "e2e_h2d": st_e["ms_h2d"]}}
    if not args.no_extras:
        ctx.pk_free(handle)
        handle = None
        single = lib.Context(0) if world > 1 else None
        line["extras"] = extras(torch, ctx, ext, dev, pk_, imad, full=args.sweep, world=world, single=single)
        bad = [k for k, v in line["extras"].items() if v.get("parity_ok") is False]
        if bad:
            raise SystemExit(f"bench.py: sweep entries fail their parity check: {bad}")
    if world == 1 and not args.no_cpu:
        # cpu_baseline: the oracle prover on the SAME instance, once, all host cores (~10-30 s); its five points double as a
        # second, Pippenger-side parity check of the GPU's
        if handle is not None:
            ctx.pk_free(handle)
        ms_cpu, pts_cpu, _ = cpu_prove(case, ocpu)
        if not np.array_equal(pts_cpu, pts_dev):
            raise SystemExit("bench.py: the CPU oracle prover and the GPU disagree on the benchmarked instance")
        line["parity"]["cpu_prover_points_equal"] = True
        line["cpu_baseline"] = {"value": ms_cpu, "unit": "ms", "cores": ocpu.threads(), "kind": "port",
                                "sample": f"one full prove of the same 2^{args.log_n}-constraint instance by the oracle/c restatement of the "
                                          "gnark-crypto CPU prover (not gnark-crypto itself), all host threads"}
    print(json.dumps(line))


def run_ours_synthetic(args):
    """BLS12-377 (BASELINE config 5) on one GPU: the pow17 MiMC-shaped chain with a trapdoor setup (scalars from the
    BLS12-377 build of oracle/c, bases from the library's fixed-base kernel), all five points checked against the
    exponent side before timing, the oracle's CPU prover on the same instance as cpu_baseline.  --synthetic-key skips the
    instance and the checks (random distinct bases with the pow17 infinity shape): only for sizes whose key the host
    cannot hold."""
    import torch
    from gnark_icicle_b200 import lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    ranks = Ranks.__new__(Ranks)
    ranks.torch, ranks.rank, ranks.world, ranks.local, ranks.dist = torch, 0, 1, 0, None
    ctx = lib.Context(0, curve=args.curve)
    ctx.set("msm_ba_rounds", args.ba_rounds)
    if args.no_tables:
        ctx.set("msm_precompute", 0)
    if args.no_g1_tables:
        ctx.set("msm_precompute_g1", 0)
    fp_eq = IMAD_EQ_PER_FP_MUL_12
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    nc = 1 << args.log_n
    case, parity, want = None, None, None
    if args.synthetic_key:
        key, sd, sh, n, nw = build_inputs(torch, ctx, nc, dev)
        handle = ctx.pk_upload(key["domain_n"], key["g1_a"], key["g1_b"], key["g1_k"], key["g1_z"], key["g2_b"], key["inf_a"], key["inf_b"], nw)
        cfg = {"workload": f"groth16_{args.curve}_prove_2p{args.log_n}_mimc17_chain", "constraints": nc, "domain": n,
               "msm_sizes": {k: int(len(key[k])) for k in ("g1_a", "g1_b", "g1_k", "g1_z", "g2_b")}, "l2": "inputs larger than L2",
               "key": "synthetic distinct bases (fixed-base kernel), MiMC pow17 chain infinity shape; no parity check on this line"}
        del key
    else:
        ocpu = oracle_for(args.curve)
        case = chain_case(nc, g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul, curve=args.curve)
        n, nw, pk = case.n, case.nw, case.pk
        cfg = workload_config(case, args.log_n, args.curve)
        handle = ctx.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, nw)
        h_cpu = ocpu.compute_h(case.a, case.b, case.c, n)
        want = case.expected_points(h_cpu)
        del h_cpu
        pin = lambda v: torch.from_numpy(np.ascontiguousarray(v).view(np.int64)).pin_memory()
        sh = {"W": pin(case.W), "a": pin(case.a), "b": pin(case.b), "c": pin(case.c)}
        sd = {"W": sh["W"].to(dev)}
        for k in ("a", "b", "c"):
            sd[k] = torch.zeros((n, 4), dtype=torch.int64, device=dev)
            sd[k][:nc] = sh[k].to(dev)
    wa, wb, wc = (torch.empty_like(sd[k]) for k in ("a", "b", "c"))
    hW, ha, hb, hc = (sh[k].numpy().view(np.uint64) for k in ("W", "a", "b", "c"))
    if case is not None:
        pts, _ = ctx.prove(handle, hW, ha, hb, hc)
        parity = parity_report(pts, want, "trapdoor identity [sum w_i s_i]G on all five points of the benchmarked instance, h from the "
                                          "CPU oracle's computeH; checked before timing", curve=args.curve)
        if parity["mismatches"]:
            raise SystemExit(f"bench.py: GPU prove disagrees with the oracle: {parity}")

    def step_dev():
        with torch.cuda.stream(ext):
            wa.copy_(sd["a"], non_blocking=True)
            wb.copy_(sd["b"], non_blocking=True)
            wc.copy_(sd["c"], non_blocking=True)
        return ctx.prove_dev(handle, sd["W"].data_ptr(), wa.data_ptr(), wb.data_ptr(), wc.data_ptr())

    l0 = lib.kernel_launch_count(args.curve)
    with ClockSampler(0) as clk:
        ms_dev, (pts_dev, st) = timed(torch, ranks, ext, step_dev, args.steps, args.warmup)
    launches = (lib.kernel_launch_count(args.curve) - l0) * args.steps // (args.steps + args.warmup)
    ms_e2e, (pts_e2e, st_e) = timed(torch, ranks, ext, lambda: ctx.prove(handle, hW, ha, hb, hc), args.steps, args.warmup)
    if not np.array_equal(pts_dev, pts_e2e):
        raise SystemExit("bench.py: device-resident and host-buffer proofs differ")
    if want is not None:
        if not (np.array_equal(pts_dev, want) and np.array_equal(pts_e2e, want)):
            raise SystemExit("bench.py: a timed prove differs from the oracle's expectation")
        parity["legs_checked"] = ["value", "e2e"]
    ctx.set("streams", 1)
    step_dev()
    _, st = step_dev()
    ctx.set("streams", 5)
    imad, wide = ctx.imad_peak()
    wide_x, fr_mul, fp_mul = ctx.pipe_rates()
    g1_ops = sum(st["msm_n"][i] * st["msm_windows"][i] for i in range(4)) * MADD_MULS["g1"] * fp_eq
    g1_ms = sum(st["ms_accumulate"][:4])
    g2_ops = st["msm_n"][4] * st["msm_windows"][4] * MADD_MULS["g2"] * fp_eq
    g2_ms = st["ms_accumulate"][4]
    ach = g1_ops / (g1_ms * 1e-3) / 1e12
    roof = {"bound": "int32-imad", "kernel": "bucket accumulation of the four G1 MSMs (batch-affine rounds + k_accumulate<G1>)", "achieved": ach,
            "peak": imad / 1e12, "unit": "TIMAD/s",
            "frac": ach / (imad / 1e12), "traffic": None, "field_mul_peak_g_per_s": fp_mul / 1e9,
            "frac_definition": "entries x 10 Fp mul x 600 IMAD-eq (12-limb field: 2*12^2+12 limb-MACs x 2) / stage time / measured IMAD peak",
            "frac_of_multiplier_ceiling": g1_ops / fp_eq / (g1_ms * 1e-3) / fp_mul,
            "g2": {"kernel": "bucket accumulation of the G2 MSM", "achieved": g2_ops / (g2_ms * 1e-3) / 1e12, "frac": g2_ops / (g2_ms * 1e-3) / imad},
            "share_of_step": (g1_ms + g2_ms) / st["ms_total"]}
    line = {"metric": f"groth16_{args.curve}_prove_ms", "value": ms_dev, "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "u32x12-montgomery",
            "data": "synthetic", "config": cfg, "parity": parity,
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": int(st_e["h2d_bytes"]), "d2h_bytes_per_step": int(st_e["d2h_bytes"])},
            "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roof, "cpu_baseline": None,
            "batch_affine_rounds": args.ba_rounds, "tables_of_multiples": "none" if args.no_tables else ("G2.B only" if args.no_g1_tables else "all five arrays"),
            "stages_ms": {"serialised_total": st["ms_total"], "compute_h": st["ms_compute_h"], "msm": st["ms_msm"],
                          "accumulate": st["ms_accumulate"], "msm_c": st["msm_c"]}}
    if case is not None and not args.no_cpu:
        ctx.pk_free(handle)
        ms_cpu, pts_cpu, _ = cpu_prove(case, ocpu)
        if not np.array_equal(pts_cpu, pts_dev):
            raise SystemExit("bench.py: the CPU oracle prover and the GPU disagree on the benchmarked instance")
        line["parity"]["cpu_prover_points_equal"] = True
        line["cpu_baseline"] = {"value": ms_cpu, "unit": "ms", "cores": ocpu.threads(), "kind": "port",
                                "sample": f"one full prove of the same 2^{args.log_n}-constraint instance by the BLS12-377 build of the oracle/c "
                                          "restatement of the gnark-crypto CPU prover (not gnark-crypto itself), all host threads"}
    print(json.dumps(line))


def extras(torch, ctx, ext, dev, pk_, imad, full=False, world=1, single=None):
    """the other two headline numbers of BASELINE.json's metric: G1 MSM Mpoints/s @ 2^24 and NTT GB/s (configs 2 and 3).
    full=True (--sweep): the whole 2^16..2^26 sweeps.  world > 1: `ctx` is a group context -- resident bases are sharded
    by point range, transforms and computeH run four-step over the GPUs with arrays on the first one (so the numbers
    include moving them over NVLink) -- and `single` is a one-GPU context used as the bit-for-bit checker.
    Every entry carries a parity flag: MSMs against the trapdoor identity (oracle/c), transforms by FFT(DIT) o
    FFTInverse(DIF) = id and (groups) equality with the single-GPU result; the oracle compares at these sizes are in
    tests/test_gpu_large.py."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import cpu as ocpu
    from helpers import g1_gen_arr
    out = {}
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)
    peak = imad * world
    hbm = pk_["hbm_gbs"] * world

    def ev_time(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            fn()
            e1.record(ext)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    host = lambda t: t.cpu().numpy().view(np.uint64)
    msm_sizes = range(16, 27) if full else ((20, 22, 24) if world == 1 else (24,))
    for logn in msm_sizes:
        m = 1 << logn
        k = rand_fr_cuda(torch, m, gen, dev)
        bases = torch.empty((m, ctx.g1w), dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        ctx.g1_fixed_base_mul_dev(k.data_ptr(), m, bases.data_ptr())
        s = rand_fr_cuda(torch, m, gen, dev)
        want = ocpu.g1_mul(g1_gen_arr(), ocpu.fr_dot(host(s), host(k)))
        if world == 1:
            got = ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m)
            ms = ev_time(lambda: ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m))
            out[f"msm_g1_2p{logn}"] = {"ms": ms, "mpts_per_s": m / ms / 1e3, "parity_ok": bool(np.array_equal(got, want))}
        # the Groth16 case: bases resident with their precomputed multiples (g16_bases_upload, one-time)
        hb = ctx.bases_upload_dev(bases.data_ptr(), m)
        del bases
        got = ctx.msm_resident_dev(hb, s.data_ptr())
        ms = ev_time(lambda: ctx.msm_resident_dev(hb, s.data_ptr()))
        windows = 255 // 20 + 1
        out[f"msm_g1_resident_2p{logn}"] = {"ms": ms, "mpts_per_s": m / ms / 1e3, "parity_ok": bool(np.array_equal(got, want)),
                                            "imad_frac_c20": m * windows * MADD_MULS["g1"] * IMAD_EQ_PER_FP_MUL / (ms * 1e-3) / peak,
                                            "gpus": world}
        ctx.bases_free(hb)
        del k, s
    if world == 1:
        m = 1 << 20
        k = rand_fr_cuda(torch, m, gen, dev)
        b2 = torch.empty((m, ctx.g2w), dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        ctx.g2_fixed_base_mul_dev(k.data_ptr(), m, b2.data_ptr())
        s = rand_fr_cuda(torch, m, gen, dev)
        ms = ev_time(lambda: ctx.msm_g2_dev(s.data_ptr(), b2.data_ptr(), m))
        out["msm_g2_2p20"] = {"ms": ms, "mpts_per_s": m / ms / 1e3}
        del k, b2, s
    ntt_sizes = range(16, 27) if full else ((20, 22, 24, 26) if world == 1 else (24, 26))
    for logn in ntt_sizes:
        m = 1 << logn
        x = rand_fr_cuda(torch, m, gen, dev)
        shapes = (("ntt_coset_dit", dict(inverse=False, coset=True, decimation=1)),)
        if full:
            shapes += (("ntt_fwd_dit", dict(inverse=False, coset=False, decimation=1)), ("intt_dif", dict(inverse=True, coset=False, decimation=0)),
                       ("intt_coset_dif", dict(inverse=True, coset=True, decimation=0)))
        for name, kw in shapes:
            ms = ev_time(lambda: ctx.ntt_dev(x.data_ptr(), m, **kw))
            gbs = 64.0 * m / (ms * 1e-3) / 1e9
            out[f"{name}_2p{logn}"] = {"ms": ms, "gb_per_s": gbs, "hbm_frac": gbs / hbm, "gpus": world,
                                       "imad_frac": (m // 2) * logn * IMAD_EQ_PER_FP_MUL / (ms * 1e-3) / peak}
        # parity: FFT(DIT) o FFTInverse(DIF) = id (different twiddle tables), and a group against the single-GPU transform
        y = x.clone()
        torch.cuda.synchronize()
        ctx.ntt_dev(y.data_ptr(), m, inverse=True, coset=False, decimation=0)
        if single is not None:
            z = x.clone()
            torch.cuda.synchronize()
            single.ntt_dev(z.data_ptr(), m, inverse=True, coset=False, decimation=0)
            torch.cuda.synchronize()
            same = bool(torch.equal(y, z))
            del z
        else:
            same = None
        ctx.ntt_dev(y.data_ptr(), m, inverse=False, coset=False, decimation=1)
        torch.cuda.synchronize()
        out[f"ntt_coset_dit_2p{logn}"]["parity_ok"] = bool(torch.equal(x, y)) and same is not False
        out[f"ntt_coset_dit_2p{logn}"]["parity"] = "round trip" + (", equals the single-GPU transform" if same else "")
        del y
        if full or world > 1 or logn in (22, 24):
            abc = torch.stack([x, rand_fr_cuda(torch, m, gen, dev), rand_fr_cuda(torch, m, gen, dev)])
            work = abc.clone()
            torch.cuda.synchronize()

            def ch():
                ctx.compute_h_dev(work[0].data_ptr(), work[1].data_ptr(), work[2].data_ptr(), m)
            ms = ev_time(ch)                      # (the inputs are whatever the previous run left: timing only)
            ent = {"ms": ms, "gb_per_s": 448.0 * m / (ms * 1e-3) / 1e9, "hbm_frac": 448.0 * m / (ms * 1e-3) / 1e9 / hbm, "gpus": world,
                   "imad_frac": 7 * (m // 2) * logn * IMAD_EQ_PER_FP_MUL / (ms * 1e-3) / peak}
            if single is not None:
                work.copy_(abc)
                ref = abc.clone()
                torch.cuda.synchronize()
                ch()
                single.compute_h_dev(ref[0].data_ptr(), ref[1].data_ptr(), ref[2].data_ptr(), m)
                torch.cuda.synchronize()
                ent["parity_ok"] = bool(torch.equal(work[0], ref[0]))
                ent["parity"] = "equals the single-GPU computeH bit for bit"
                del ref
            out[f"compute_h_2p{logn}"] = ent
            del abc, work
        del x
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-n", type=int, default=22, help="log2 of the constraint count (BASELINE: 22)")
    ap.add_argument("--cpu-budget-s", type=float, default=300.0, help="--impl reference: stop timing further proves once the timed region exceeds this")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg of the GPU arm")
    ap.add_argument("--ba-rounds", type=int, default=4, help="batch-affine rounds of the MSMs (library default 4; 0 = XYZZ accumulation only)")
    ap.add_argument("--curve", default="bn254", choices=["bn254", "bls12_377"],
                    help="bls12_377 = BASELINE config 5 (use --log-n 24); the default line is the BN254 headline")
    ap.add_argument("--g2-weight", type=int, default=0, help="multi-GPU key placement: cost of a G2 point in hundredths of a G1 point (0 = library default by group size)")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--synthetic-key", action="store_true", help="bls12_377: random bases with the pow17 infinity shape instead of the trapdoor instance (no parity, no CPU leg)")
    ap.add_argument("--no-g1-tables", action="store_true", help="bls12_377: a table of multiples for G2.B only")
    ap.add_argument("--no-tables", action="store_true", help="bls12_377: keep no tables of multiples (memory goes to the batch-affine buffers instead)")
    ap.add_argument("--sweep", action="store_true", help="extras = the full 2^16..2^26 MSM and NTT sweeps (configs 2, 3)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
