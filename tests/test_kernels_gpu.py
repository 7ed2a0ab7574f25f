"""Parity of every C-ABI entry point against the Tier-S oracle (oracle/w8a8.py) on seeded inputs.

Bar: int8 activations, row scales, int32 accumulators and fp16 GEMM outputs BIT-EXACT; attention
and lm_head (fp accumulation order differs) within the tolerance written in each test.
"""
import numpy as np
import pytest
import torch

import b200stage
from b200stage import ops
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def n(x):
    return x.detach().cpu().numpy()


def rng(seed):
    return np.random.default_rng(seed)


# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,d", [(1, 256), (7, 2048), (300, 4096), (33, 5120), (5, 768)])
@pytest.mark.parametrize("ln", [False, True])
@pytest.mark.parametrize("smooth", [False, True])
def test_norm_quant_exact(T, d, ln, smooth):
    r = rng(T * 131 + d)
    x = (r.standard_normal((T, d)) * r.uniform(0.1, 4.0, (T, 1))).astype(np.float16)
    x[0, : min(d, 16)] *= 30           # outliers
    g = (1 + 0.1 * r.standard_normal(d)).astype(np.float16)
    b = (0.05 * r.standard_normal(d)).astype(np.float16) if ln else None
    sm = r.uniform(0.3, 3.0, d).astype(np.float32) if smooth else None
    q_ref, s_ref = O.norm_quant(x, g, b, sm, 1e-5)
    q, s = ops.norm_quant(t(x), t(g), None if b is None else t(b), None if sm is None else t(sm), 1e-5)
    assert np.array_equal(n(s), s_ref)
    assert np.array_equal(n(q), q_ref)


def test_norm_quant_zero_row_and_empty():
    d = 512
    x = np.zeros((3, d), np.float16)
    x[1] = 1.0
    g = np.ones(d, np.float16)
    q, s = ops.norm_quant(t(x), t(g), None, None, 1e-5)
    q_ref, s_ref = O.norm_quant(x, g, None, None, 1e-5)
    assert np.array_equal(n(q), q_ref) and np.array_equal(n(s), s_ref)
    assert n(s)[0] == 0 and not n(q)[0].any()
    q, s = ops.norm_quant(t(np.zeros((0, d), np.float16)), t(g), None, None, 1e-5)
    assert q.shape == (0, d)


@pytest.mark.parametrize("T,d", [(4, 2048), (65, 4096)])
def test_final_norm_exact(T, d):
    r = rng(d + T)
    x = r.standard_normal((T, d)).astype(np.float16)
    g = (1 + 0.1 * r.standard_normal(d)).astype(np.float16)
    y = ops.norm(t(x), t(g), None, 1e-5)
    assert np.array_equal(n(y), O.final_norm(x, g, None, 1e-5))


@pytest.mark.parametrize("T,K", [(1, 4096), (9, 2048), (130, 5120)])
def test_smooth_quant_exact(T, K):
    r = rng(K + T)
    x = (r.standard_normal((T, K)) * 2).astype(np.float16)
    sm = r.uniform(0.2, 5.0, K).astype(np.float32)
    q, s = ops.smooth_quant(t(x), t(sm))
    q_ref, s_ref = O.smooth_quant(x, sm)
    assert np.array_equal(n(s), s_ref) and np.array_equal(n(q), q_ref)


@pytest.mark.parametrize("T,F", [(1, 5632), (6, 11008), (70, 14336), (3, 1024)])
def test_act_mul_quant_exact(T, F):
    r = rng(F + T)
    gu = (r.standard_normal((T, 2 * F)) * 3).astype(np.float16)
    gu[0, :8] = [0, -0.0, 60, -60, 11.5, -11.5, 1e-3, -1e-3]
    sm = r.uniform(0.2, 5.0, F).astype(np.float32)
    q, s = ops.act_mul_quant(t(gu), t(sm), ops.ACT_SILU_MUL)
    q_ref, s_ref = O.act_mul_quant(gu, sm, "silu")
    assert np.array_equal(n(s), s_ref) and np.array_equal(n(q), q_ref)


def test_act_gelu_erf_quant_exact():
    """exact-erf GELU (GPT-NeoX / Pythia `hidden_act = "gelu"`), ungated, with smoothing."""
    r = rng(6)
    h = (r.standard_normal((9, 8192)) * 3).astype(np.float16)
    h[0, :6] = [0, -0.0, 9, -9, 0.5, -0.5]
    sm = r.uniform(0.2, 5.0, 8192).astype(np.float32)
    q, s = ops.act_mul_quant(t(h), t(sm), ops.ACT_GELU_ERF)
    q_ref, s_ref = O.act_quant(h, sm, "gelu")
    assert np.array_equal(n(s), s_ref) and np.array_equal(n(q), q_ref)


def test_act_gelu_quant_exact():
    r = rng(5)
    h = (r.standard_normal((9, 3072)) * 3).astype(np.float16)
    q, s = ops.act_mul_quant(t(h), None, ops.ACT_GELU)
    q_ref, s_ref = O.act_quant(h, None, "gelu_new")
    assert np.array_equal(n(s), s_ref) and np.array_equal(n(q), q_ref)


# ----------------------------------------------------------------------------------------------
def _gemm_case(M, N, K, bias, resid, sched, seed=0):
    r = rng(seed + M * 7 + N * 3 + K)
    a = r.integers(-127, 128, (M, K), dtype=np.int8)
    w = r.integers(-127, 128, (N, K), dtype=np.int8)
    sa = r.uniform(1e-3, 5e-2, M).astype(np.float32)
    swt = r.uniform(1e-4, 2e-3, N).astype(np.float32)
    b = (0.1 * r.standard_normal(N)).astype(np.float16) if bias else None
    res = r.standard_normal((M, N)).astype(np.float16) if resid else None
    out_ref, acc_ref = O.gemm_w8a8(a, w, sa, swt, b, res)
    acc = torch.full((M, N), -7, dtype=torch.int32, device=DEV)
    out = ops.gemm_w8a8(t(a), t(w), t(sa), t(swt), None if b is None else t(b), None if res is None else t(res),
                        acc_dump=acc, schedule=sched)
    torch.cuda.synchronize()
    assert np.array_equal(n(acc), acc_ref), f"int32 accumulators differ: {np.abs(n(acc).astype(np.int64) - acc_ref).max()}"
    assert np.array_equal(n(out).view(np.uint16), out_ref.view(np.uint16))


@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (128, 128, 512), (300, 512, 1024), (130, 2560, 2048),
                                   (512, 4096, 4096), (200, 64, 256), (77, 40, 144), (2048, 1024, 2048),
                                   (16, 4096, 4096), (64, 11008, 256)])
def test_gemm_tc_exact(M, N, K):
    _gemm_case(M, N, K, bias=(N % 512 == 0), resid=(M % 2 == 0), sched=1)


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("N,K", [(4096, 4096), (2560, 2048), (1000, 11008), (24, 144)])
def test_gemm_stream_exact(M, N, K):
    _gemm_case(M, N, K, bias=True, resid=True, sched=2)


def test_gemm_auto_and_schedules_agree():
    for M in (1, 8, 9, 129):
        _gemm_case(M, 512, 1024, True, True, 0, seed=3)
    _gemm_case(4, 512, 1024, True, True, 1, seed=3)      # tensor-core path on a skinny M


def test_gemm_extremes_exact():
    """|acc| at its maximum (all +-127, K = 14336): int32 headroom and float(acc) rounding."""
    M, N, K = 130, 256, 14336
    a = np.full((M, K), 127, np.int8)
    w = np.full((N, K), -127, np.int8)
    w[::2] = 127
    sa = np.full(M, 1e-2, np.float32)
    swt = np.full(N, 1e-3, np.float32)
    out_ref, acc_ref = O.gemm_w8a8(a, w, sa, swt)
    acc = torch.zeros((M, N), dtype=torch.int32, device=DEV)
    out = ops.gemm_w8a8(t(a), t(w), t(sa), t(swt), acc_dump=acc, schedule=1)
    assert np.array_equal(n(acc), acc_ref) and np.array_equal(n(out).view(np.uint16), out_ref.view(np.uint16))


def test_gemm_linearity_full_size():
    """Size-independent property at a BASELINE-sized shape (Llama-2-7B gate_up, 2048 tokens):
    acc(A, W1 + W2) == acc(A, W1) + acc(A, W2) exactly, and row 0 equals the streaming schedule's."""
    M, N, K = 2048, 22016, 4096
    g = torch.Generator(device=DEV).manual_seed(1)
    a = torch.randint(-127, 128, (M, K), generator=g, device=DEV, dtype=torch.int8)
    w1 = torch.randint(-63, 64, (N, K), generator=g, device=DEV, dtype=torch.int8)
    w2 = torch.randint(-63, 64, (N, K), generator=g, device=DEV, dtype=torch.int8)
    one_m = torch.ones(M, device=DEV)
    one_n = torch.ones(N, device=DEV)
    accs = []
    for w in (w1, w2, w1 + w2):
        acc = torch.empty((M, N), dtype=torch.int32, device=DEV)
        ops.gemm_w8a8(a, w.contiguous(), one_m, one_n, acc_dump=acc, schedule=1)
        accs.append(acc)
    assert torch.equal(accs[0] + accs[1], accs[2])
    acc1 = torch.empty((1, N), dtype=torch.int32, device=DEV)
    ops.gemm_w8a8(a[:1].contiguous(), w1, one_m[:1], one_n, acc_dump=acc1, schedule=2)
    assert torch.equal(acc1[0], accs[0][0])


def test_gemm_rejects_bad_shapes():
    a = torch.zeros((4, 100), dtype=torch.int8, device=DEV)      # K % 16 != 0
    w = torch.zeros((8, 100), dtype=torch.int8, device=DEV)
    s = torch.ones(8, device=DEV)
    with pytest.raises(b200stage._abi.B2Error):
        ops.gemm_w8a8(a, w, s[:4], s)


# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("H,KV,hd", [(4, 4, 64), (8, 2, 64), (4, 4, 128), (32, 8, 128)])
def test_rope_kv_append_exact(H, KV, hd):
    r = rng(H + KV + hd)
    T, Smax, rows = 37, 128, 3
    qkv = r.standard_normal((T, (H + 2 * KV) * hd)).astype(np.float16)
    pos = r.permutation(Smax)[:T].astype(np.int32)
    seq = r.integers(0, rows, T).astype(np.int32)
    cos, sin = O.rope_table(hd, Smax)
    kc = torch.zeros((rows, KV, Smax, hd), dtype=torch.float16, device=DEV)
    vc = torch.zeros_like(kc)
    d_qkv = t(qkv)
    ops.rope_kv_append(d_qkv, t(cos), t(sin), t(pos), kc, vc, H, KV, hd, seq_ids=t(seq))
    q_ref = O.apply_rope(qkv[:, :H * hd].reshape(T, H, hd), pos, cos, sin)
    k_ref = O.apply_rope(qkv[:, H * hd:(H + KV) * hd].reshape(T, KV, hd), pos, cos, sin)
    v_ref = qkv[:, (H + KV) * hd:].reshape(T, KV, hd)
    got = n(d_qkv)
    assert np.array_equal(got[:, :H * hd].reshape(T, H, hd), q_ref)
    kcn, vcn = n(kc), n(vc)
    for i in range(T):
        assert np.array_equal(kcn[seq[i], :, pos[i]], k_ref[i])
        assert np.array_equal(vcn[seq[i], :, pos[i]], v_ref[i])
    # the table the package builds is the oracle's table
    c2, s2 = b200stage.weights.rope_table(hd, Smax)
    assert np.array_equal(c2.numpy(), cos) and np.array_equal(s2.numpy(), sin)


def _attn_tol(got, ref):
    """Tolerance for attention outputs (fp16, |o| <= max|v|): max-abs <= 2e-3 * max|ref| + 1e-3,
    cosine >= 0.99999."""
    got = got.astype(np.float64).ravel()
    ref = ref.astype(np.float64).ravel()
    err = np.abs(got - ref).max()
    cos = float(got @ ref / (np.linalg.norm(got) * np.linalg.norm(ref) + 1e-30))
    assert err <= 2e-3 * np.abs(ref).max() + 1e-3, err
    assert cos >= 0.99999, cos


@pytest.mark.parametrize("H,KV,hd,B,S,window", [
    (4, 4, 64, 2, 70, 0), (8, 2, 64, 3, 300, 0), (32, 32, 128, 1, 1500, 0), (32, 8, 128, 2, 2100, 0),
    (32, 4, 64, 1, 513, 0), (8, 2, 128, 2, 700, 256), (4, 4, 64, 1, 1, 0), (40, 40, 128, 4, 33, 0),
    (32, 8, 128, 1, 161, 96), (32, 8, 128, 1, 8000, 4096), (32, 32, 96, 2, 1300, 0), (4, 4, 96, 1, 70, 0), (8, 8, 256, 2, 1100, 0), (2, 2, 256, 1, 40, 0)])
def test_decode_attn(H, KV, hd, B, S, window):
    r = rng(H * 3 + S)
    Smax = S + 8
    kc = r.standard_normal((B, KV, Smax, hd)).astype(np.float16)
    vc = r.standard_normal((B, KV, Smax, hd)).astype(np.float16)
    q = r.standard_normal((B, (H + 2 * KV) * hd)).astype(np.float16)     # q inside a qkv-shaped buffer
    seqlens = np.array([max(1, S - 17 * b) for b in range(B)], np.int32)
    scale = 1.0 / np.sqrt(hd)
    nbytes = ops.decode_attn_workspace_bytes(B, H, KV, hd, int(seqlens.max()), window)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=DEV)
    out = ops.decode_attn(t(q), t(kc), t(vc), t(seqlens), H, KV, hd, int(seqlens.max()), scale, window, workspace=ws)
    for b in range(B):
        L = int(seqlens[b])
        ref = O.attention(q[b:b + 1, :H * hd].reshape(1, H, hd), np.transpose(kc[b, :, :L], (1, 0, 2)),
                          np.transpose(vc[b, :, :L], (1, 0, 2)), np.array([L - 1]), scale, window)
        _attn_tol(n(out)[b].reshape(H, hd), ref[0])
    # without a workspace the kernel must still be right (fewer splits)
    out2 = ops.decode_attn(t(q), t(kc), t(vc), t(seqlens), H, KV, hd, int(seqlens.max()), scale, window)
    _attn_tol(n(out2), n(out))


@pytest.mark.parametrize("H,KV,hd,T,pos0,window", [
    (4, 4, 64, 64, 0, 0), (4, 2, 64, 100, 0, 0), (8, 8, 128, 130, 0, 0), (4, 4, 64, 50, 77, 0),
    (8, 2, 128, 200, 56, 0), (4, 2, 64, 150, 30, 48), (2, 2, 64, 1, 0, 0), (2, 1, 128, 9, 300, 100), (4, 4, 96, 200, 17, 0),
    (32, 32, 96, 300, 0, 0), (8, 8, 256, 200, 33, 0), (2, 2, 256, 70, 0, 0)])
def test_prefill_attn(H, KV, hd, T, pos0, window):
    r = rng(H + T + pos0)
    S = pos0 + T
    Smax = S + 5
    kc = r.standard_normal((2, KV, Smax, hd)).astype(np.float16)
    vc = r.standard_normal((2, KV, Smax, hd)).astype(np.float16)
    qkv = r.standard_normal((T, (H + 2 * KV) * hd)).astype(np.float16)
    scale = 1.0 / np.sqrt(hd)
    out = ops.prefill_attn(t(qkv), t(kc), t(vc), pos0, 1, H, KV, hd, scale, window)
    ref = O.attention(qkv[:, :H * hd].reshape(T, H, hd), np.transpose(kc[1, :, :S], (1, 0, 2)),
                      np.transpose(vc[1, :, :S], (1, 0, 2)), np.arange(pos0, S), scale, window)
    _attn_tol(n(out).reshape(T, H, hd), ref)


# ----------------------------------------------------------------------------------------------
def test_embed_lm_head_argmax():
    r = rng(11)
    V, d = 1000, 512
    table = (0.16 * r.standard_normal((V, d))).astype(np.float16)
    ids = r.integers(0, V, 19).astype(np.int64)
    e = ops.embed(t(ids), t(table))
    assert np.array_equal(n(e), table[ids])
    for M in (1, 5, 19, 140):
        x = r.standard_normal((M, d)).astype(np.float16)
        logits = ops.lm_head(t(x), t(table))
        ref = O.lm_head(x, table)
        # fp32 accumulation in a different order: |err| <= 1e-3 * max|ref|
        assert np.abs(n(logits) - ref).max() <= 1e-3 * np.abs(ref).max()
        am = ops.argmax(logits)
        assert np.array_equal(n(am), n(logits).argmax(-1))


@pytest.mark.parametrize("M,V,d", [(1, 32000, 4096), (1, 16000, 4096), (2, 50257, 768), (3, 1000, 256), (4, 32001, 2048), (1, 7, 64)])
@pytest.mark.parametrize("merge", [False, True])
def test_lm_head_argmax_one_launch(M, V, d, merge):
    """b2_lm_head_argmax = b2_lm_head + b2_argmax_shard, exactly: same logits, same winner (ties to the lowest
    index), same merge with another shard's winner; the scratch comes back zeroed, so launches can follow each
    other on one stream."""
    r = rng(M * 7 + V + d + merge)
    x = t((r.standard_normal((M, d)) * 0.5).astype(np.float16))
    Wn = (r.standard_normal((V, d)) * 0.05).astype(np.float16)
    if V > 100:                                     # exact ties: the same row at several indices, two of them the maximum
        Wn[V - 3] = Wn[5]
        Wn[40] = Wn[5]
        big = (n(x)[0].astype(np.float32) * 0.1).astype(np.float16)
        Wn[V // 2] = big
        Wn[V // 2 - 9] = big
    W = t(Wn)
    off = 3000 if merge else 0
    scratch = ops.head_argmax_scratch(DEV)
    logits_ref = ops.lm_head(x, W)
    other = None
    if merge:
        ov = n(logits_ref).max(-1).astype(np.float32).copy()
        oi = np.arange(M, dtype=np.int64) + 17
        if M > 1:
            ov[1] = np.float32(-1e9)                # this row's other shard loses
        if M > 2:
            oi[2] = off + V + 100                   # equal value, higher index: ours wins the tie
        other = (t(ov), t(oi))                      # (row 0: equal value, lower index -> the other shard wins)
    idx_ref = torch.empty((M,), dtype=torch.int64, device=DEV)
    val_ref = torch.empty((M,), dtype=torch.float32, device=DEV)
    ops.argmax_shard(logits_ref, off, out_idx=idx_ref, out_val=val_ref, other=other)
    for rep in range(3):
        logits = torch.full((M, V), -7.0, dtype=torch.float32, device=DEV) if rep == 0 else None
        idx2 = torch.full((M,), -1, dtype=torch.int64, device=DEV)
        val = torch.empty((M,), dtype=torch.float32, device=DEV)
        idx = ops.lm_head_argmax(x, W, scratch, idx_offset=off, other=other, out_idx2=idx2, out_val=val, logits=logits)
        assert np.array_equal(n(idx), n(idx_ref)), (rep, n(idx), n(idx_ref))
        assert np.array_equal(n(idx2), n(idx_ref))
        assert np.array_equal(n(val), n(val_ref))
        if logits is not None:
            assert np.array_equal(n(logits), n(logits_ref))
        assert int(scratch.count_nonzero()) == 0
    if V > 100 and not merge:
        assert int(n(idx_ref)[0]) == V // 2 - 9     # the tie resolved to the lower of the two maximal rows


def test_shared_head_protocol_on_one_gpu():
    """The three launches of the head shared by all pipeline stages, with every "peer" on this GPU: the hidden state
    goes to W inboxes at once (b2_p2p_send_multi), each slice's b2_lm_head_argmax waits for its copy (the flag is
    already there: nothing spins), stores its winner record and publishes it, b2_merge_winners picks the token.
    Result = b2_argmax over the whole row."""
    r = rng(77)
    W, B, V, d = 5, 3, 4096 + 40, 512
    x = t((r.standard_normal((B, d)) * 0.5).astype(np.float16))
    Wn = (r.standard_normal((V, d)) * 0.05).astype(np.float16)
    Wn[3000] = Wn[10]                                   # equal logits in two different slices: the lower index wins
    big = (n(x)[1].astype(np.float32) * 0.1).astype(np.float16)
    Wn[10], Wn[3000] = big, big
    head = t(Wn)
    want = n(ops.argmax(ops.lm_head(x, head)))
    assert want[1] == 10
    cuts = [((V * k // W) & ~7) for k in range(W)] + [V]
    inbox = torch.zeros((W, B, d), dtype=torch.float16, device=DEV)
    nflag = torch.zeros((W,), dtype=torch.int32, device=DEV)
    records = torch.zeros((W * 64,), dtype=torch.uint8, device=DEV)
    wflag = torch.zeros((W,), dtype=torch.int32, device=DEV)
    err = torch.zeros((1,), dtype=torch.int32, device=DEV)
    scratch = ops.head_argmax_scratch(DEV)
    md = ops.MultiDst([(inbox[k], nflag[k: k + 1]) for k in range(W)])
    for seq in (1, 2):
        ops.p2p_send_multi(x, md, seq)
        for k in range(W):
            idx = records[k * 64: k * 64 + 32].view(torch.int64)[:B]
            val = records[k * 64 + 32: k * 64 + 48].view(torch.float32)[:B]
            ops.lm_head_argmax(inbox[k], head[cuts[k]: cuts[k + 1]].contiguous(), scratch, idx_offset=cuts[k], out_idx=idx, out_val=val,
                               wait=(nflag[k: k + 1], seq), notify=(wflag[k: k + 1], seq), err=err)
        tok = torch.full((B,), -1, dtype=torch.int64, device=DEV)
        tok2 = torch.full((B,), -1, dtype=torch.int64, device=DEV)
        ops.merge_winners(wflag, seq, records, W, B, tok, tok2, err)
        assert np.array_equal(n(tok), want) and np.array_equal(n(tok2), want), (seq, n(tok), want)
        assert np.array_equal(n(nflag), np.full(W, seq)) and np.array_equal(n(wflag), np.full(W, seq))
        assert np.array_equal(n(inbox), np.broadcast_to(n(x), (W, B, d)))
    assert int(err.item()) == 0 and int(scratch.count_nonzero()) == 0


def test_lm_head_argmax_refuses_wide_batches():
    x = torch.zeros((5, 64), dtype=torch.float16, device=DEV)
    W = torch.zeros((32, 64), dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError):
        ops.lm_head_argmax(x, W, ops.head_argmax_scratch(DEV))


@pytest.mark.parametrize("V", [1001, 50257])
def test_lm_head_argmax_sample_any_vocabulary_size(V):
    """GPT-2's vocabulary (50257) is odd: the logits row pitch is then not a multiple of 16 bytes."""
    r = rng(13)
    d = 64
    table = (0.16 * r.standard_normal((V, d))).astype(np.float16)
    for M in (1, 3, 8, 9, 140):
        x = r.standard_normal((M, d)).astype(np.float16)
        logits = ops.lm_head(t(x), t(table))
        ref = O.lm_head(x, table)
        assert logits.shape == (M, V) and np.abs(n(logits) - ref).max() <= 1e-3 * np.abs(ref).max()
        assert np.array_equal(n(ops.argmax(logits)), n(logits).argmax(-1))
    got = ops.sample(logits[:3], temperature=0.9, top_k=40, top_p=0.95, seed=3, step=7)     # rows at an odd pitch
    lg = n(logits)
    assert [int(x) for x in got] == [O.sample_det(lg[b], np.zeros((0,), np.int64), 0.9, 40, 0.95, 1.0, 3, 7, b) for b in range(3)]


def test_embed_with_positions_exact():
    r = rng(12)
    V, d, P = 300, 768, 64
    table = (0.1 * r.standard_normal((V, d))).astype(np.float16)
    ptab = (0.02 * r.standard_normal((P, d))).astype(np.float16)
    ids = r.integers(0, V, 10).astype(np.int64)
    pos = r.integers(0, P, 10).astype(np.int32)
    e = ops.embed(t(ids), t(table), t(pos), t(ptab))
    ref = (table[ids].astype(np.float32) + ptab[pos].astype(np.float32)).astype(np.float16)
    assert np.array_equal(n(e), ref)
