"""computeH across GPUs: one process (and one libg16b200 context) per GPU, the four-step split of every transform
with one all-to-all per transform (SURVEY.md 8(e) level 3).  torch / NCCL is the plumbing for the exchange only.

Layout: a, b, c enter CYCLIC-distributed (rank r holds x[r::world], zero padded to n/world); the transforms
alternate between the cyclic and the block distribution (g16_dist_ntt_stages_dev); h leaves BLOCK-distributed in
bit-reversed order, i.e. rank r holds h[r n/world : (r+1) n/world] -- exactly the range its Z-base shard covers
(g16_pk_upload with shard_index/shard_count), so the H.Z MSM needs no further data movement.

`compute_h_steps` is a generator: it yields (send, recv) wire tensors at each of the three exchange points and the
driver performs the all-to-all (NCCL in `compute_h_distributed`; an in-process shuffle in the single-GPU tests).
"""
import torch

TW_FWD, TW_INV, TW_COSET_FWD, TW_COSET_INV = 0, 1, 2, 3


def _log2(v):
    l = v.bit_length() - 1
    assert 1 << l == v, f"{v} is not a power of two"
    return l


def split_stage(n, world):
    """global stage index where the transforms switch distribution: any value in [g, L - g] works."""
    L, g = _log2(n), _log2(world)
    assert L >= 2 * g, "domain too small for this many ranks"
    return max(g, min(L - g, L // 2))


def cyclic_slice(x, n, rank, world):
    """numpy/torch array of n_constraints rows -> this rank's zero-padded cyclic part (n/world rows)."""
    import numpy as np
    n_local = n // world
    part = x[rank::world]
    out = np.zeros((n_local,) + tuple(x.shape[1:]), dtype=x.dtype)
    out[: len(part)] = part
    return out


def compute_h_steps(ctx, abc, n, rank, world):
    """abc: int64 CUDA tensor (3, n/world, 4) = cyclic parts of a, b, c (clobbered).  Generator; returns via
    StopIteration.value the (n/world, 4) view holding this rank's block of h."""
    n_local = n // world
    X = split_stage(n, world)
    wire_s = torch.empty_like(abc)
    wire_r = torch.empty_like(abc)
    p, ps, pr = abc.data_ptr(), wire_s.data_ptr(), wire_r.data_ptr()

    def exchange(count, from_block):
        ctx.dist_permute_dev(p, ps, count, n_local, world, block_layout=int(from_block), to_wire=1)
        return wire_s[:count].view(world, -1), wire_r[:count].view(world, -1)

    def land(count, to_block):
        ctx.dist_permute_dev(pr, p, count, n_local, world, block_layout=int(to_block), to_wire=0)

    # 1. three inverse transforms (DIF): top stages cyclic, all-to-all, low stages block
    ctx.dist_ntt_stages_dev(p, 3, n, world, rank, TW_INV, dif=1, cyclic=1, split=X)
    yield exchange(3, from_block=False)
    land(3, to_block=True)
    ctx.dist_ntt_stages_dev(p, 3, n, world, rank, TW_INV, dif=1, cyclic=0, split=X)
    # 2. three coset transforms (DIT): low stages block, all-to-all, top stages cyclic
    ctx.dist_ntt_stages_dev(p, 3, n, world, rank, TW_COSET_FWD, dif=0, cyclic=0, split=X)
    yield exchange(3, from_block=True)
    land(3, to_block=False)
    ctx.dist_ntt_stages_dev(p, 3, n, world, rank, TW_COSET_FWD, dif=0, cyclic=1, split=X)
    # 3. pointwise quotient on the cyclic parts, result in a
    ctx.dist_pointwise_h_dev(p, n_local, n)
    # 4. inverse coset transform of a (DIF): cyclic, all-to-all, block
    ctx.dist_ntt_stages_dev(p, 1, n, world, rank, TW_COSET_INV, dif=1, cyclic=1, split=X)
    yield exchange(1, from_block=False)
    land(1, to_block=True)
    ctx.dist_ntt_stages_dev(p, 1, n, world, rank, TW_COSET_INV, dif=1, cyclic=0, split=X)
    return abc[0]


def compute_h_distributed(ctx, abc, n, rank, world, group=None, stream=None):
    """NCCL driver of compute_h_steps (torch.distributed must be initialised when world > 1).

    stream: a torch stream wrapping g16_ctx_stream(ctx) (torch.cuda.ExternalStream).  When given, the whole pipeline is
    stream-ordered: the library calls only enqueue ("dist_sync" 0), the all-to-alls are issued with that stream current
    (NCCL orders itself after the work already on it and the stream waits for NCCL), and the host never waits -- the MSM
    call that consumes h runs on the same stream.  Without it every step is followed by a host synchronisation."""
    import torch.distributed as dist
    if stream is not None:
        ctx.set("dist_sync", 0)
        try:
            with torch.cuda.stream(stream):
                gen = compute_h_steps(ctx, abc, n, rank, world)
                try:
                    while True:
                        send, recv = next(gen)
                        if world > 1:
                            dist.all_to_all_single(recv, send, group=group)
                        else:
                            recv.copy_(send)
                except StopIteration as done:
                    return done.value
        finally:
            ctx.set("dist_sync", 1)
    gen = compute_h_steps(ctx, abc, n, rank, world)
    try:
        while True:
            send, recv = next(gen)
            if world > 1:
                dist.all_to_all_single(recv, send, group=group)
                torch.cuda.current_stream().synchronize()
            else:
                recv.copy_(send)
                torch.cuda.current_stream().synchronize()
    except StopIteration as done:
        return done.value


def compute_h_emulated(ctx, parts, n):
    """single-GPU emulation of `world = len(parts)` ranks (tests): steps every rank's generator in lockstep and
    performs the all-to-all as device copies.  parts[r] = (3, n/world, 4) cyclic tensors; returns the blocks of h."""
    world = len(parts)
    gens = [compute_h_steps(ctx, parts[r], n, r, world) for r in range(world)]
    out = [None] * world
    while True:
        wires = []
        for r, g in enumerate(gens):
            try:
                wires.append(next(g))
            except StopIteration as done:
                out[r] = done.value
        if not wires:
            return out
        assert len(wires) == world
        for src in range(world):
            for dst in range(world):
                wires[dst][1][src].copy_(wires[src][0][dst])
        torch.cuda.synchronize()
