#!/usr/bin/env python
"""bench.py -- BN254 Groth16 prove @ 2^22 constraints on N B200s (BASELINE.json configs[3]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--log-n 22] [--no-extras]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the proving hot path (computeH + the five MSMs + window Horner on the host) over one
synthetic R1CS solution.  `value` is timed with the inputs resident in HBM (g16_prove_bn254_dev); `e2e` goes
through the plugin call with pinned HOST buffers (g16_prove_bn254: H2D of W,a,b,c inside the timed region, the
five result points read back).  The proving key is a DummySetup-style synthetic key (distinct random bases
[k_i]G made by the library's fixed-base kernel; the reference's own BenchmarkProver also uses a dummy key,
backend/groth16/groth16_test.go:78) with the MiMC-chain shape |A| = |K| ~ n, |B| ~ 2n/3, |Z| = n-1.
Only the cpu_baseline / --impl reference legs touch oracle/ (the CPU restatement of the reference prover).
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


def chain_matrices(nc, power, coeffs):
    """constraint matrices (CSR) with the MiMC-chain shape for the wires-only entry point: per
    round t = x + k + c_i, then the squarings, then v * t; `coeffs` = [one, c_1, c_2, ...] (random constants)."""
    per = 3 if power == 5 else 5
    rounds = (nc - 1) // per
    r = np.arange(rounds, dtype=np.int64)
    base = 4 + per * r                                   # first new wire of round r
    xw = np.where(r == 0, 2, base - 1)
    t_w = np.stack([xw, np.full(rounds, 3), np.zeros(rounds, dtype=np.int64)], axis=1)
    t_c = np.stack([np.zeros(rounds, dtype=np.int64), np.zeros(rounds, dtype=np.int64), 1 + r % (len(coeffs) - 1)], axis=1)
    z1 = np.zeros((rounds, 1), dtype=np.int64)
    col = lambda a: a.reshape(-1, 1)
    sq = [(col(base + i), z1) for i in range(per - 1)]   # u, v (, v2, v3): squared by the next row, the last times t
    Lr = [(t_w, t_c)] + sq
    Rr = [(t_w, t_c)] + sq[:-1] + [(t_w, t_c)]
    Or = [(col(base + i), z1) for i in range(per)]
    nfill = nc - 1 - per * rounds
    f = np.arange(nfill, dtype=np.int64)
    last_state = int(base[-1] + per - 1) if rounds else 2
    fx_in = np.where(f == 0, last_state, 3 + per * rounds + f)
    fx_out = 4 + per * rounds + f
    final_state = int(fx_out[-1]) if nfill else last_state

    def matrix(per_round, fill_w, last_w):
        lens = [p[0].shape[1] for p in per_round]
        w = np.concatenate([p[0] for p in per_round], axis=1).reshape(-1)
        c = np.concatenate([p[1] for p in per_round], axis=1).reshape(-1)
        row_len = np.concatenate([np.tile(lens, rounds), np.ones(nfill, dtype=np.int64), [len(last_w)]])
        rp = np.zeros(nc + 1, dtype=np.uint32)
        rp[1:] = np.cumsum(row_len)
        return (rp, np.concatenate([w, fill_w, last_w]).astype(np.uint32),
                np.concatenate([c, np.zeros(nfill + len(last_w), dtype=np.int64)]).astype(np.uint32))

    return matrix(Lr, fx_in, [final_state, 3]), matrix(Rr, fx_in, [0]), matrix(Or, fx_out, [1])


def one_fr(curve):
    """Montgomery image of 1 (R mod r) as 4 uint64 limbs."""
    r = {"bn254": 21888242871839275222246405745257275088548364400416034343698204186575808495617,
         "bls12_377": 8444461749428370424248824938781546531375899335154063827935233455917409239041}[curve]
    v = (1 << 256) % r
    return np.array([(v >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)


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


# ------------------------------------------------------------------------------------------------ CPU leg (oracle)
def cpu_prove_sample(log_nc, steps=1, warmup=0):
    """The reference's CPU prover algorithm (backend/groth16/bn254/prove.go:141-320) as restated in oracle/c,
    timed on the host cores on a bounded sample: a full prove at 2^log_nc constraints.  Returns ms per prove."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import cpu as ocpu
    from helpers import ChainCase
    nc = 1 << log_nc
    case = ChainCase(nc, seed=11)
    pk = case.pk
    wa, wb, wk = case.W[~case.inf_a], case.W[~case.inf_b], case.W[2:]
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        h = ocpu.compute_h(case.a, case.b, case.c, case.n)
        ocpu.msm_g1(pk.g1_a, wa)
        ocpu.msm_g1(pk.g1_b, wb)
        ocpu.msm_g1(pk.g1_k, wk)
        ocpu.msm_g1(pk.g1_z, h[: case.n - 1])
        ocpu.msm_g2(pk.g2_b, wb)
        if it >= warmup:
            times.append((time.perf_counter() - t0) * 1e3)
    return sum(times) / len(times), ocpu.threads()


def cpu_scale(log_s, log_t):
    """bucket-method work ~ points x windows; the oracle picks c = log2(n) - 4 (<= 16) like gnark-crypto's size table."""
    c = lambda l: min(16, max(4, l - 4))
    return 2.0 ** (log_t - log_s) * (c(log_s) / c(log_t))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    log_s = args.cpu_log_n
    ms, cores = cpu_prove_sample(log_s, steps=max(1, args.steps), warmup=min(args.warmup, 1))
    scale = cpu_scale(log_s, args.log_n)
    v = ms * scale
    sample = (f"CPU restatement (oracle/c, not gnark-crypto: no Go toolchain) of a full prove at 2^{log_s} constraints, "
              f"x{scale:.1f} extrapolation to 2^{args.log_n} (points x windows)")
    line = {"impl": "reference", "metric": "groth16_bn254_prove_ms", "value": v, "unit": "ms", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": v, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32x8-montgomery", "data": "synthetic",
            "config": {"workload": f"groth16_bn254_prove_2p{args.log_n}_mimc_chain", "constraints": 1 << args.log_n},
            "cpu_baseline": {"value": v, "unit": "ms", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ GPU arm
def timed(torch, ext, fn, steps, warmup, world, dist):
    """W warm-up calls, then K calls bracketed by barrier + synchronize and CUDA events on the library's stream."""
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    last = None
    for _ in range(steps):
        last = fn()
    e1.record(ext)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps, last


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gnark_icicle_b200 import lib
    import gnark_icicle_b200 as g16

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the accelerated path has no CPU fallback")
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    if world > 1:
        # NCCL prints its version banner on stdout when the communicator comes up (NCCL_DEBUG=VERSION from the box's
        # environment or nccl.conf); rank 0 must print ONE JSON line there, so stdout points at stderr until then
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    ctx = lib.Context(local, curve=args.curve)
    fp_eq = IMAD_EQ_PER_FP_MUL if args.curve == "bn254" else IMAD_EQ_PER_FP_MUL_12
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    nc = 1 << args.log_n
    key, sd, sh, n, nw = build_inputs(torch, ctx, nc, dev)
    handle = ctx.pk_upload(key["domain_n"], key["g1_a"], key["g1_b"], key["g1_k"], key["g1_z"], key["g2_b"],
                           key["inf_a"], key["inf_b"], nw, shard_index=rank, shard_count=world)
    sizes = {k: int(len(key[k])) for k in ("g1_a", "g1_b", "g1_k", "g1_z", "g2_b")}
    del key
    wa, wb, wc = (torch.empty_like(sd[k]) for k in ("a", "b", "c"))
    gather = [torch.empty(ctx.prove_out_words, dtype=torch.int64, device=dev) for _ in range(world)] if world > 1 else None

    def combine(pts):
        if world == 1:
            return pts
        t = torch.from_numpy(pts.view(np.int64)).to(dev)
        dist.all_gather(gather, t)                      # 384 B per rank over NCCL/NVLink
        if rank != 0:
            return pts
        return g16.combine_shards([g.cpu().numpy().view(np.uint64) for g in gather], curve=args.curve)

    hW, ha, hb, hc = (sh[k].numpy().view(np.uint64) for k in ("W", "a", "b", "c"))
    if world == 1:
        def step_dev():
            with torch.cuda.stream(ext):                    # a,b,c are clobbered by the transform pipeline
                wa.copy_(sd["a"], non_blocking=True)
                wb.copy_(sd["b"], non_blocking=True)
                wc.copy_(sd["c"], non_blocking=True)
            pts, st = ctx.prove_dev(handle, sd["W"].data_ptr(), wa.data_ptr(), wb.data_ptr(), wc.data_ptr())
            return combine(pts), st

        def step_e2e():
            pts, st = ctx.prove(handle, hW, ha, hb, hc)     # pinned host buffers in, 5 points out
            return combine(pts), st
    else:
        # four-step computeH across the ranks (gnark_icicle_b200/dist.py): every rank owns the cyclic part of a, b, c,
        # ends with its block of h, and runs the five MSMs over its point-range shard of the key
        from gnark_icicle_b200 import dist as gdist
        n_local = n // world
        abc_master = torch.stack([sd[k][rank::world] for k in ("a", "b", "c")]).contiguous()
        abc_work = torch.empty_like(abc_master)
        abc_pin = abc_master.cpu().pin_memory()
        # G16_DIST_ASYNC=1: stream-ordered distributed computeH (no host waits between its steps).  Measured at N=2:
        # 40.9 vs 40.5 ms host-synchronised -- the GPU is the bottleneck, not the host -- so the plain form is the default
        dist_stream = ext if os.environ.get("G16_DIST_ASYNC") else None
        W_stage = torch.zeros_like(sd["W"])
        w_lo, w_hi = ctx.pk_wire_range(handle)           # the wires this rank's shard of the A, B, K bases reads
        del wa, wb, wc

        def msms(W_t):
            torch.cuda.current_stream().synchronize()
            # (g16_prove_msms_begin_dev could start the W MSMs here, under the distributed computeH; measured: 41.3 vs 40.4 ms
            #  at N=2, 13.7 vs 13.8 at N=8 -- the GPUs are busy either way, and h arriving late leaves the H.Z MSM alone at the end)
            h_blk = gdist.compute_h_distributed(ctx, abc_work, n, rank, world, stream=dist_stream)
            pts, st = ctx.prove_msms_dev(handle, W_t.data_ptr(), h_blk.data_ptr())
            return combine(pts), st

        def step_dev():
            abc_work.copy_(abc_master, non_blocking=True)
            return msms(sd["W"])

        def step_e2e():
            # this rank's 1/world of a, b, c and the wire range its key shard reads, from pinned host memory
            abc_work.copy_(abc_pin, non_blocking=True)
            W_stage[w_lo:w_hi].copy_(sh["W"][w_lo:w_hi], non_blocking=True)
            pts, st = msms(W_stage)
            st["h2d_bytes"] = (w_hi - w_lo) * W_stage.shape[1] * 8 + abc_pin.numel() * 8
            return pts, st

    if world > 1:
        # one-time check of the multi-rank path: this rank's block of the distributed h == the same block of h computed
        # by this GPU alone from the full a, b, c
        abc_work.copy_(abc_master)
        torch.cuda.synchronize()
        h_blk = gdist.compute_h_distributed(ctx, abc_work, n, rank, world, stream=dist_stream)
        torch.cuda.synchronize()
        full = [sd[k].clone() for k in ("a", "b", "c")]
        torch.cuda.synchronize()                     # the clones were made on torch's stream, computeH runs on the library's
        ctx.compute_h_dev(full[0].data_ptr(), full[1].data_ptr(), full[2].data_ptr(), n)
        torch.cuda.synchronize()
        if not torch.equal(h_blk, full[0][rank * n_local:(rank + 1) * n_local]):
            raise SystemExit(f"bench.py: rank {rank}: distributed computeH differs from the single-GPU h")
        del full
    l0 = lib.kernel_launch_count(args.curve)
    with ClockSampler(local) as clk:
        ms_dev, (pts_dev, st) = timed(torch, ext, step_dev, args.steps, args.warmup, world, dist)
    launches = (lib.kernel_launch_count(args.curve) - l0) * args.steps // (args.steps + args.warmup)
    ms_e2e, (pts_e2e, st_e) = timed(torch, ext, step_e2e, args.steps, args.warmup, world, dist)
    if rank == 0 and not np.array_equal(pts_dev, pts_e2e):
        raise SystemExit("bench.py: device-resident and host-buffer proofs differ")
    e2e_wires = None
    if world == 1:
        # constraint matrices resident, only the wire vector crosses PCIe
        gen = torch.Generator(device=dev)
        gen.manual_seed(99)
        coeffs = np.concatenate([ctx.fr_mul_vec(np.zeros((0, 4), dtype=np.uint64), np.zeros((0, 4), dtype=np.uint64)),
                                 rand_fr_cuda(torch, 1025, gen, dev).cpu().numpy().view(np.uint64)])
        coeffs[0] = one_fr(args.curve)
        r1 = ctx.r1cs_upload(nc, nw, coeffs, chain_matrices(nc, 5 if args.curve == "bn254" else 17, coeffs))
        ms_w, (_, st_w) = timed(torch, ext, lambda: ctx.prove_wires(handle, r1, hW), args.steps, args.warmup, world, dist)
        e2e_wires = {"value": ms_w, "unit": "ms", "h2d_bytes_per_step": int(st_w["h2d_bytes"]), "d2h_bytes_per_step": int(st_w["d2h_bytes"]),
                     "note": "g16_prove_bn254_wires: a, b, c evaluated on the device from resident constraint matrices (8(f) N1)"}
        ctx.r1cs_free(r1)

    # per-kernel durations for the roofline: one extra step with the MSMs serialised (with overlap on, kernels of
    # different MSMs share the SMs and their individual event times are not meaningful)
    ctx.set("streams", 1)
    step_dev()
    _, st = step_dev()
    ctx.set("streams", 5)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    imad, wide = ctx.imad_peak()
    wide_x, fr_mul, fp_mul = ctx.pipe_rates()
    # dominant kernel: k_accumulate<G1> (4 launches per step); algorithmic IMAD-equivalents per launch =
    # entries (points x windows) x 10 Fp multiplications per XYZZ mixed add x 272
    g1_ops = sum(st["msm_n"][i] * st["msm_windows"][i] for i in range(4)) * MADD_MULS["g1"] * fp_eq
    g1_ms = sum(st["ms_accumulate"][:4])
    g2_ops = st["msm_n"][4] * st["msm_windows"][4] * MADD_MULS["g2"] * fp_eq
    g2_ms = st["ms_accumulate"][4]
    ach = g1_ops / (g1_ms * 1e-3) / 1e12
    roof = {"bound": "int32-imad", "kernel": "k_accumulate<G1> (4 launches/step)", "achieved": ach, "peak": imad / 1e12,
            "unit": "TIMAD/s", "frac": ach / (imad / 1e12), "traffic": None,
            "peak_source": "g16_imad_peak microbenchmark, same run (independent mad.lo.u32 chains on all SMs)",
            "imad_wide_peak_tops": wide / 1e12, "imad_wide_carry_peak_tops": wide_x / 1e12,
            "field_mul_peak_g_per_s": fp_mul / 1e9, "field_mul_achieved_g_per_s": g1_ops / fp_eq / (g1_ms * 1e-3) / 1e9,
            "frac_of_multiplier_ceiling": g1_ops / fp_eq / (g1_ms * 1e-3) / fp_mul,
            "g2": {"kernel": "k_accumulate<G2>", "achieved": g2_ops / (g2_ms * 1e-3) / 1e12,
                   "frac": g2_ops / (g2_ms * 1e-3) / imad},
            "share_of_step": (g1_ms + g2_ms) / st["ms_total"],
            "note": "kernel times from one extra step with the five MSMs serialised (streams=1); the timed steps overlap them"}
    pk_, src = peaks()
    # HBM view of the same kernel: 64 B affine point + 4 B sorted index per bucket entry; `traffic` is
    # dram__bytes_read+write per launch of the 2^22-point launch from the ncu --set full capture in profiles/
    g1_bytes = sum(st["msm_n"][i] * st["msm_windows"][i] for i in range(4)) * (8 * ctx.g1w + 4)
    roof["hbm"] = {"bound": "hbm", "achieved": g1_bytes / (g1_ms * 1e-3) / 1e9, "peak": pk_["hbm_gbs"], "unit": "GB/s",
                   "frac": g1_bytes / (g1_ms * 1e-3) / 1e9 / pk_["hbm_gbs"], "peak_source": src,
                   "algorithmic_bytes_per_launch_2p22": (1 << 22) * 13 * 68,
                   "note": "the accumulate kernel is bound by the integer multiplier, not HBM: it moves <10 % of the copy bandwidth"}
    if args.log_n == 22 and args.curve == "bn254":
        roof["traffic"] = 4.10e9
        roof["traffic_source"] = ("profiles/r01d_accumulate_ntt_2p22_ncu_raw.csv: 3.97 GB read + 0.13 GB written per 2^22-point launch "
                                  "(8.6 ms, 54.5 M entries x 68 B = 3.71 GB algorithmic); before the 64-byte L2 fill hint: 7.45 GB")
    ntt_bytes = 7 * 64 * n
    if world == 1 and st["ms_compute_h"] > 0:
      roof["compute_h"] = {"bound": "hbm", "achieved": ntt_bytes / (st["ms_compute_h"] * 1e-3) / 1e9, "peak": pk_["hbm_gbs"],
                         "unit": "GB/s", "frac": ntt_bytes / (st["ms_compute_h"] * 1e-3) / 1e9 / pk_["hbm_gbs"],
                         "peak_source": src, "note": "7 transforms x 64n B algorithmic; the butterflies are IMAD-bound at 254 bits"}
    else:
      roof["compute_h"] = {"note": "distributed four-step computeH: timed inside the step, not per kernel"}
    cname, shape = args.curve, ("mimc_chain" if args.curve == "bn254" else "mimc17_chain")
    line = {"metric": f"groth16_{cname}_prove_ms", "value": ms_dev, "unit": "ms", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32x8-montgomery" if cname == "bn254" else "u32x12-montgomery", "data": "synthetic",
            "config": {"workload": f"groth16_{cname}_prove_2p{args.log_n}_{shape}", "constraints": nc, "domain": n,
                       "msm_sizes": sizes, "parallelism": f"MSM point-range shards x{world}" + (", four-step computeH (NCCL all-to-all)" if world > 1 else ""), "l2": "inputs larger than L2",
                       "key": "synthetic distinct bases (fixed-base kernel), MiMC-chain infinity shape"},
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": int(st_e["h2d_bytes"]),
                    "d2h_bytes_per_step": int(st_e["d2h_bytes"])},
            "e2e_wires": e2e_wires, "gpu_launches": int(launches), "clocks": clk.summary(), "roofline": roof,
            "stages_ms": {"serialised_total": st["ms_total"], "compute_h": st["ms_compute_h"], "msm": st["ms_msm"], "accumulate": st["ms_accumulate"],
                          "host_epilogue": st["ms_host"], "msm_c": st["msm_c"], "e2e_h2d": st_e["ms_h2d"]}}
    if world == 1 and args.curve != "bn254":
        line["cpu_baseline"] = None
        line["cpu_baseline_note"] = ("the C restatement of the CPU prover (oracle/c) exists for BN254 only; the BLS12-377 "
                                     "oracle is Python big-int code and is a checker, not a timing baseline")
        if not args.no_extras:
            line["extras"] = extras(torch, ctx, ext, dev, pk_, imad, full=args.sweep)
    elif world == 1 and not args.no_extras:
        line["extras"] = extras(torch, ctx, ext, dev, pk_, imad, full=args.sweep)
        ms_cpu, cores = cpu_prove_sample(args.cpu_log_n)
        scale = cpu_scale(args.cpu_log_n, args.log_n)
        line["cpu_baseline"] = {"value": ms_cpu * scale, "unit": "ms", "cores": cores, "kind": "port",
                                "sample": f"oracle/c CPU restatement of the reference prover (not gnark-crypto), full prove at "
                                          f"2^{args.cpu_log_n} constraints = {ms_cpu:.0f} ms, x{scale:.1f} extrapolation (points x windows)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def extras(torch, ctx, ext, dev, pk_, imad, full=False):
    """the other two headline numbers of BASELINE.json's metric: G1 MSM Mpoints/s @ 2^24 and NTT GB/s.
    full=True (--sweep): the whole 2^16..2^26 sweeps of configs 2 and 3."""
    out = {}
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)

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

    for logn in (range(16, 27) if full else (20, 22, 24)):
        m = 1 << logn
        k = rand_fr_cuda(torch, m, gen, dev)
        bases = torch.empty((m, ctx.g1w), dtype=torch.int64, device=dev)
        ctx.g1_fixed_base_mul_dev(k.data_ptr(), m, bases.data_ptr())
        s = rand_fr_cuda(torch, m, gen, dev)
        ms = ev_time(lambda: ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m))
        out[f"msm_g1_2p{logn}"] = {"ms": ms, "mpts_per_s": m / ms / 1e3}
        # the Groth16 case: bases resident with their precomputed multiples (g16_bases_upload, one-time)
        hb = ctx.bases_upload_dev(bases.data_ptr(), m)
        del bases
        ms = ev_time(lambda: ctx.msm_resident_dev(hb, s.data_ptr()))
        out[f"msm_g1_resident_2p{logn}"] = {"ms": ms, "mpts_per_s": m / ms / 1e3}
        ctx.bases_free(hb)
        del k, s
    m = 1 << 20
    k = rand_fr_cuda(torch, m, gen, dev)
    b2 = torch.empty((m, ctx.g2w), dtype=torch.int64, device=dev)
    ctx.g2_fixed_base_mul_dev(k.data_ptr(), m, b2.data_ptr())
    s = rand_fr_cuda(torch, m, gen, dev)
    ms = ev_time(lambda: ctx.msm_g2_dev(s.data_ptr(), b2.data_ptr(), m))
    out["msm_g2_2p20"] = {"ms": ms, "mpts_per_s": m / ms / 1e3}
    del k, b2, s
    for logn in (range(16, 27) if full else (20, 22, 24, 26)):
        m = 1 << logn
        x = rand_fr_cuda(torch, m, gen, dev)
        if full:
            for name, kw in (("ntt_fwd_dit", dict(inverse=False, coset=False, decimation=1)),
                             ("intt_dif", dict(inverse=True, coset=False, decimation=0)),
                             ("intt_coset_dif", dict(inverse=True, coset=True, decimation=0))):
                ms = ev_time(lambda: ctx.ntt_dev(x.data_ptr(), m, **kw))
                out[f"{name}_2p{logn}"] = {"ms": ms, "gb_per_s": 64.0 * m / (ms * 1e-3) / 1e9}
            abc = [rand_fr_cuda(torch, m, gen, dev) for _ in range(3)]
            ms = ev_time(lambda: ctx.compute_h_dev(abc[0].data_ptr(), abc[1].data_ptr(), abc[2].data_ptr(), m))
            out[f"compute_h_2p{logn}"] = {"ms": ms, "gb_per_s": 448.0 * m / (ms * 1e-3) / 1e9}
            del abc
        ms = ev_time(lambda: ctx.ntt_dev(x.data_ptr(), m, inverse=False, coset=True, decimation=1))
        gbs = 64.0 * m / (ms * 1e-3) / 1e9
        muls = (m // 2) * logn
        out[f"ntt_coset_dit_2p{logn}"] = {"ms": ms, "gb_per_s": gbs, "hbm_frac": gbs / pk_["hbm_gbs"],
                                         "imad_frac": muls * IMAD_EQ_PER_FP_MUL / (ms * 1e-3) / imad}
        del x
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-n", type=int, default=22, help="log2 of the constraint count (BASELINE: 22)")
    ap.add_argument("--cpu-log-n", type=int, default=18, help="log2 constraints of the bounded CPU sample")
    ap.add_argument("--curve", default="bn254", choices=["bn254", "bls12_377"],
                    help="bls12_377 = BASELINE config 5 (use --log-n 24); the default line is the BN254 headline")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="extras = the full 2^16..2^26 MSM and NTT sweeps (configs 2, 3)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "ours" and args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
