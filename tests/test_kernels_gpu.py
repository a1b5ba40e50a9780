"""CUDA kernels vs plain PyTorch fp32 references / the bit-exact host mirrors.  Run on the B200 box."""
import ctypes

import pytest
import torch

from biscotti_b200.ops import ml, krum as krum_ops, secagg
from biscotti_b200.ops.msm import CommitmentEngine
from biscotti_b200.crypto import primitives as cp

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_native_library_is_loaded_and_counts_launches(lib):
    lib.bsc_launch_count_reset()
    ml.noise_row(64, 1, 2, 3, 1.0, device=DEV)
    torch.cuda.synchronize()
    assert lib.bsc_launch_count() == 1


@pytest.mark.parametrize("P,B,F,C", [(7, 10, 784, 10), (3, 10, 20, 4), (2, 5, 8742, 2)])
def test_softmax_step_matches_fp32_reference(P, B, F, C):
    torch.manual_seed(0)
    w = (torch.randn(C * F + C) * 0.05)
    X = torch.randn(P, 32, F)
    y = torch.randint(0, C, (P, 32), dtype=torch.int32)
    idx = torch.stack([torch.randperm(32)[:B] for _ in range(P)]).int()
    Xb = torch.gather(X, 1, idx.long().unsqueeze(-1).expand(P, B, F))
    yb = torch.gather(y, 1, idx.long())
    Uref, lref = ml.softmax_step_reference(w, Xb, yb, C, 100.0)
    U, loss = ml.softmax_step(w.to(DEV), X.to(DEV), y.to(DEV), idx.to(DEV), C, 100.0)
    # (10 classes, 784 features take the tensor-core kernel: logits from bf16 hi/lo operand planes, ~2^-16 relative)
    assert torch.allclose(U.cpu(), Uref, rtol=2e-4, atol=2e-6)
    assert torch.allclose(loss.cpu(), lref, rtol=1e-4, atol=1e-5)
    # clipping active: scale gradients up through tiny clip
    Uref2, _ = ml.softmax_step_reference(w, Xb, yb, C, 0.01)
    U2, _ = ml.softmax_step(w.to(DEV), X.to(DEV), y.to(DEV), idx.to(DEV), C, 0.01)
    assert torch.allclose(U2.cpu(), Uref2, rtol=1e-4, atol=1e-7)
    assert abs(float(U2[0].norm()) - 0.01) < 1e-4


@pytest.mark.parametrize("P,B", [(100, 10), (200, 10), (301, 10), (150, 16), (4, 32), (1, 3)])
def test_softmax_step_tensor_core_kernel_matches_simt_kernel_and_reference(lib, P, B):
    """K1 on tcgen05 (logits GEMM in TMEM, bf16x3 operands) against the SIMT kernel and the fp32 PyTorch restatement;
    one, two and three peers per CTA (by peer count), the last CTA partly filled."""
    torch.manual_seed(5)
    F, C, Ns = 784, 10, 40
    w = torch.randn(C * F + C) * 0.05
    X = torch.randn(P, Ns, F)
    y = torch.randint(0, C, (P, Ns), dtype=torch.int32)
    idx = torch.stack([torch.randperm(Ns)[:B] for _ in range(P)]).int()
    Xb = torch.gather(X, 1, idx.long().unsqueeze(-1).expand(P, B, F))
    yb = torch.gather(y, 1, idx.long())
    Uref, lref = ml.softmax_step_reference(w, Xb, yb, C, 100.0)
    args = (w.to(DEV), X.to(DEV), y.to(DEV), idx.to(DEV), C, 100.0)
    lib.bsc_softmax_force_simt(1)
    try:
        Us, ls = ml.softmax_step(*args)
        torch.cuda.synchronize()
    finally:
        lib.bsc_softmax_force_simt(0)
    Ut, lt = ml.softmax_step(*args, ldu=C * F + C + 2)            # padded row stride: the pad must come back zero
    torch.cuda.synchronize()
    assert torch.allclose(Us.cpu(), Uref, rtol=1e-4, atol=1e-6)
    assert torch.allclose(Ut[:, : C * F + C].cpu(), Uref, rtol=2e-4, atol=2e-6)
    assert torch.allclose(lt.cpu(), lref, rtol=1e-4, atol=1e-5)
    assert float(Ut[:, C * F + C:].abs().max()) == 0.0


def test_logistic_step_matches_reference():
    torch.manual_seed(1)
    P, B, F = 9, 10, 25
    w = torch.randn(F) * 0.1
    X = torch.randn(P, B, F)
    y = torch.sign(torch.randn(P, B))
    ref = ml.logistic_step_reference(w, X, y, 1e-2, 1e-2)
    out = ml.logistic_step(w.to(DEV), X.to(DEV), y.to(DEV), None, 1e-2, 1e-2)
    assert torch.allclose(out.cpu()[:, :F], ref, rtol=1e-4, atol=1e-7)


def test_noise_rows_match_host_generator_and_are_gaussian():
    d = 7850
    g = ml.noise_row(d, 42, 5, 17, 1.0, device=DEV).cpu()
    h = ml.noise_row(d, 42, 5, 17, 1.0, device="cpu")
    assert torch.allclose(g, h, rtol=1e-4, atol=1e-4)
    big = torch.cat([ml.noise_row(d, 42, p, 0, 1.0, device=DEV) for p in range(20)]).cpu()
    assert abs(float(big.mean())) < 0.02 and abs(float(big.std()) - 1.0) < 0.02
    assert not torch.equal(ml.noise_row(d, 42, 5, 18, 1.0, device=DEV).cpu(), g)


def test_noise_mix_planes_and_norms():
    torch.manual_seed(2)
    P, d = 6, 1000
    ldu = 1000
    U = torch.randn(P, ldu) * 0.1
    noisers = torch.tensor([[1, 2], [0, -1], [-1, -1], [3, 4], [5, 0], [2, 2]], dtype=torch.int32)
    ref = ml.noise_mix(U, noisers, 9, 3, -0.5, d)
    out = ml.noise_mix(U.to(DEV), noisers.to(DEV), 9, 3, -0.5, d)
    assert torch.allclose(out[0].cpu(), ref[0], rtol=1e-4, atol=1e-5)
    assert torch.allclose(out[3].cpu(), ref[3], rtol=1e-4, atol=1e-4)
    rec = out[1].float() + out[2].float()
    assert torch.allclose(rec[:, :d].cpu(), ref[0][:, :d], rtol=2e-5, atol=1e-6)     # bf16 hi+lo ~ 16 mantissa bits
    assert float(out[1][:, d:].float().abs().sum()) == 0.0
    assert torch.equal(out[0][2].cpu(), U[2])                                      # no noisers => untouched


@pytest.mark.parametrize("n,d,ksplit", [(70, 7850, 8), (33, 257, 1), (140, 1000, 3)])
def test_gram_simt_and_krum_select_match_reference(n, d, ksplit):
    torch.manual_seed(3)
    X = torch.randn(n, d) * 0.1
    X[: n // 5] += 0.5                      # a poisoned cluster
    Xd = X.to(DEV)
    G = krum_ops.gram_simt(Xd, d, ksplit).sum(0).cpu()
    assert torch.allclose(G, X @ X.t(), rtol=1e-4, atol=1e-3)
    clip = n // 2
    sref, aref, Dref = krum_ops.krum_reference(X, clip)
    s, a, D = krum_ops.krum_select(krum_ops.gram_simt(Xd, d, ksplit), (Xd * Xd).sum(1), n, clip, want_D=True)
    assert torch.allclose(D.cpu().double(), Dref, rtol=1e-3, atol=1e-2)
    assert torch.allclose(s.cpu().double(), sref, rtol=1e-3, atol=1e-1)
    assert a.cpu().bool().tolist() == aref.tolist()
    assert int(a.sum()) == n - clip


def test_secagg_kernels_match_host_mirrors():
    torch.manual_seed(4)
    P, d, S = 12, 7850, 21
    U = torch.randn(P, 7852) * 0.05
    Ud = U.to(DEV)
    Q = secagg.quantize(Ud, d)
    assert torch.equal(Q.cpu(), secagg.quantize(U, d))
    Y = secagg.shares_eval(Q, S)
    assert torch.equal(Y.cpu(), secagg.shares_eval(Q.cpu(), S))
    sel = torch.tensor([0, 3, 4, 9, 11], dtype=torch.int32)
    Ysum = secagg.share_sum(Ud, sel.to(DEV), d, S)
    assert torch.equal(Ysum.cpu(), Y.cpu()[sel.long()].sum(0))
    # reconstruct from non-contiguous shares (miner 1 of 3 missing) and apply the step
    xs = torch.tensor([0, 1, 2, 3, 4, 5, 6, 14, 15, 16], dtype=torch.int32, device=DEV)
    avail = torch.tensor([1] * 7 + [0] * 7 + [1] * 7, dtype=torch.uint8, device=DEV)
    w64 = torch.zeros(d, dtype=torch.float64, device=DEV)
    w32 = torch.zeros(d, device=DEV)
    dsum, bad = secagg.reconstruct_step(Ysum, xs, avail, d, w64, w32, want_dsum=True)
    assert int(bad.sum()) == 0
    exp = Q.cpu()[sel.long()][:, :d].sum(0).double() / 1e4
    assert torch.equal(w64.cpu(), exp) and torch.equal(dsum.cpu(), exp)
    assert torch.equal(w32.cpu(), exp.float())
    Ybad = Ysum.clone()
    Ybad[5, 20] += 3
    _, bad = secagg.reconstruct_step(Ybad, xs, avail, d, torch.zeros_like(w64), None)
    assert int(bad[1]) == 1
    # plain aggregation
    w = torch.zeros(d, dtype=torch.float64, device=DEV)
    secagg.plain_sum_step(Ud, sel.to(DEV), d, w, None)
    assert torch.allclose(w.cpu(), U[sel.long()][:, :d].double().sum(0))


def test_msm_kernels_match_host_mirror_bit_for_bit():
    d, S = 57, 21
    cpu = CommitmentEngine(d, S, tau=2, device="cpu")
    gpu = CommitmentEngine(d, S, tau=2, device=DEV)
    assert gpu.to_bytes(gpu.wtable) == cpu.to_bytes(cpu.wtable)
    g = torch.Generator().manual_seed(5)
    Q = torch.zeros(5, 60, dtype=torch.int64)
    Q[:, :d] = torch.randint(-10**6, 10**6, (5, d), generator=g)
    Q[1] = 0                                              # all-zero update -> infinity everywhere
    Q[2, 3] = -(2**40)                                    # a large negative scalar
    Cc, Cg = cpu.chunk_commitments(Q), gpu.chunk_commitments(Q.to(DEV))
    assert gpu.to_bytes(Cg) == cpu.to_bytes(Cc)
    assert gpu.to_bytes(gpu.full_commitments(Cg)) == cpu.to_bytes(cpu.full_commitments(Cc))
    Wc, Wg = cpu.witnesses(Q), gpu.witnesses(Q.to(DEV))
    assert gpu.to_bytes(Wg) == cpu.to_bytes(Wc)
    sel = torch.tensor([0, 2, 4], dtype=torch.int32)
    assert gpu.to_bytes(gpu.aggregate(Wg, sel.to(DEV))) == cpu.to_bytes(cpu.aggregate(Wc, sel))
    # random-tau SRS generated on the device equals the host one
    cpu_r = CommitmentEngine(20, S, tau=987654321, device="cpu")
    gpu_r = CommitmentEngine(20, S, tau=987654321, device=DEV)
    assert cpu_r.srs_bytes == gpu_r.srs_bytes


def test_batched_vrf_and_sha256_on_device(lib):
    import hashlib

    keys = [cp.VRFKey(hashlib.sha256(bytes([i])).digest()) for i in range(9)]
    msg = hashlib.sha256(b"blk").digest()
    assert cp.BatchedVRF(keys, DEV).outputs(msg) == [k.prove(msg)[0] for k in keys]
    msgs = torch.randint(0, 256, (50, 100), dtype=torch.uint8)
    out = torch.empty(50, 32, dtype=torch.uint8, device=DEV)
    md = msgs.to(DEV)
    rc = lib.bsc_sha256_batch(ctypes.c_void_p(md.data_ptr()), 100, 100, 50, ctypes.c_void_p(out.data_ptr()),
                              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    for i in (0, 17, 49):
        assert bytes(out[i].cpu().tolist()) == hashlib.sha256(bytes(msgs[i].tolist())).digest()


def test_eval_kernels():
    torch.manual_seed(6)
    N, F, C, M = 300, 784, 10, 4
    X = torch.randn(N, F)
    y = torch.randint(0, C, (N,), dtype=torch.int32)
    Wc = torch.randn(M, C * F + C) * 0.1
    base = torch.randn(C * F + C) * 0.1
    ref = ml.eval_softmax_errors(Wc, base, X, y, C)
    out = ml.eval_softmax_errors(Wc.to(DEV), base.to(DEV), X.to(DEV), y.to(DEV), C)
    assert torch.allclose(out.cpu(), ref, atol=2.0 / N)
    Xl = torch.randn(N, 25)
    yl = torch.sign(torch.randn(N))
    Wl = torch.randn(M, 25)
    assert torch.allclose(ml.eval_logistic_errors(Wl.to(DEV), None, Xl.to(DEV), yl.to(DEV)).cpu(),
                          ml.eval_logistic_errors(Wl, None, Xl, yl), atol=2.0 / N)


@pytest.mark.parametrize("n,d,ksplit", [(100, 7850, 16), (70, 7850, None), (128, 64, 1), (200, 1000, 4), (300, 2500, 5)])
def test_gram_tcgen05_matches_fp32_reference(n, d, ksplit):
    """tcgen05/TMA Gram kernel (bf16x3 split, fp32 TMEM accumulation) vs an fp64 reference of the same
    product; the dropped lo*lo term bounds the relative error by ~2^-16."""
    from biscotti_b200.ops import gram_tc

    torch.manual_seed(7)
    U = torch.randn(n, ml.pad_to(d, 4)) * 0.1
    U[:, d:] = 0
    noisers = torch.full((n, 1), -1, dtype=torch.int32)
    Un, hi, lo, norms = ml.noise_mix(U.to(DEV), noisers.to(DEV), 1, 0, 0.0, d)
    if ksplit is None:
        ksplit = gram_tc.ksplit_for(n, hi.shape[1])
    Gp = gram_tc.gram_tc(hi, lo, n, ksplit)
    torch.cuda.synchronize()
    G = Gp.sum(0)[:n, :n].double().cpu()
    X = U[:, :d].double()
    ref = X @ X.t()
    scale = float(ref.diagonal().mean())
    assert torch.allclose(G, ref, rtol=0, atol=3e-5 * scale), float((G - ref).abs().max() / scale)
    assert torch.allclose(G, G.t(), rtol=0, atol=3e-5 * scale)  # off-diagonal tiles are mirrored copies
    # end to end through krum(): same accept set as the fp64 restatement
    clip = n // 2
    _, aref, _ = krum_ops.krum_reference(U[:, :d], clip)
    _, a, _ = krum_ops.krum(Un, d, clip, hi, lo, norms, impl="tcgen05")
    assert a.cpu().bool().tolist() == aref.tolist()


def test_vrf_side_stream_path_matches_generic_lottery():
    """Engine path: radix-2^25.5 Ed25519 kernel on a side stream + one native noiser-lottery call, against
    the generic-field prove() outputs and the per-peer Python-driven lottery."""
    import hashlib

    import numpy as np

    n = 40
    keys = [cp.VRFKey(hashlib.sha256(b"k%d" % i).digest()) for i in range(n)]
    msg = hashlib.sha256(b"prev-block").digest()
    stake = [10 + 5 * (i % 4) for i in range(n)]
    bv = cp.BatchedVRF(keys, DEV)
    is_worker = np.ones(n, dtype=np.uint8)
    is_worker[[3, 7]] = 0
    tab = bv.noiser_table(bv.launch(msg), stake, is_worker, 2)
    for p in range(n):
        if not is_worker[p]:
            assert tab[p].tolist() == [-1, -1]
            continue
        out = keys[p].prove(msg)[0]
        assert tab[p].tolist() == cp.lottery_noisers(out, stake, p, 2)


def test_vrf_quad_and_single_thread_kernels_agree():
    import hashlib

    import numpy as np

    n = 37
    keys = [cp.VRFKey(hashlib.sha256(b"q%d" % i).digest()) for i in range(n)]
    msg = hashlib.sha256(b"another-block").digest()
    stake = [10] * n
    w = np.ones(n, dtype=np.uint8)
    quad = cp.BatchedVRF(keys, DEV, kernel="quad")
    one = cp.BatchedVRF(keys, DEV, kernel="r25")
    hq, ho = quad.launch(msg), one.launch(msg)
    tq = quad.noiser_table(hq, stake, w, 3)
    to = one.noiser_table(ho, stake, w, 3)
    assert bytes(hq[0].numpy().tobytes()) == bytes(ho[0].numpy().tobytes())     # Gamma_p = x_p * H, all peers
    assert bytes(hq[0][5].tolist()) == keys[5].prove(msg)[1][64:]
    assert (tq == to).all()


def test_batched_schnorr_on_device_matches_host_and_verifies():
    import hashlib
    import struct

    from biscotti_b200.crypto.schnorr_batch import SchnorrBatchSigner

    secrets = [hashlib.sha256(b"sk%d" % i).digest() for i in range(7)]
    tags = [struct.pack("<qii", 3, i, 2 * i) for i in range(7)]
    msgs = [bytes([i]) * 64 for i in range(7)]
    gpu, cpu = SchnorrBatchSigner(DEV), SchnorrBatchSigner("cpu")
    sg = gpu.finish(gpu.launch(secrets, tags), msgs)
    sc = cpu.finish(cpu.launch(secrets, tags), msgs)
    assert sg == sc                                           # comb kernel T = v*G equals the host double-and-add
    for sk, m, s in zip(secrets, msgs, sg):
        assert cp.schnorr_verify(m, cp.schnorr_pubkey(sk), s)


@pytest.mark.parametrize("kind", ["mnist_cnn", "cifar_cnn", "lfw_cnn"])
def test_cnn_local_step_kernels_match_autograd(kind):
    """Hand-written conv/fc forward+backward (kernels_cnn.cu) vs per-peer torch autograd on the CPU."""
    from biscotti_b200.models.registry import CNN_ARCH, cnn_local_step, cnn_param_slices
    from biscotti_b200.ops import cnn_native

    arch = CNN_ARCH[kind]
    _, d = cnn_param_slices(arch)
    torch.manual_seed(8)
    P, B = 3, 10
    F = arch["image"][0] * arch["image"][1] * arch["image"][2]
    w = torch.randn(d) * 0.02
    Xb = torch.randn(P, B, F)
    yb = torch.randint(0, arch["n_classes"], (P, B), dtype=torch.int32)
    ref = cnn_local_step(arch, w, Xb, yb, 100.0, torch.zeros(P, d))
    out = torch.zeros(P, d + 2, device=DEV)
    cnn_native.local_step(kind, w.to(DEV), Xb.to(DEV), yb.to(DEV), 100.0, out)
    got = out.cpu()[:, :d]
    scale = float(ref.abs().max())
    assert torch.allclose(got, ref, rtol=1e-3, atol=2e-5 * max(1.0, scale)), float((got - ref).abs().max())
    # clipping path
    ref2 = cnn_local_step(arch, w, Xb, yb, 0.05, torch.zeros(P, d))
    cnn_native.local_step(kind, w.to(DEV), Xb.to(DEV), yb.to(DEV), 0.05, out)
    assert torch.allclose(out.cpu()[:, :d], ref2, rtol=1e-3, atol=1e-6)
    assert abs(float(out[0, :d].norm()) - 0.05) < 1e-3


@pytest.mark.parametrize("kind", ["mnist_cnn", "cifar_cnn"])
def test_cnn_shape_specialised_kernels_match_generic_kernels(lib, kind):
    """The register-resident kernels for the two reference one-conv shapes against the generic ones, at a size that
    spans several image groups of the classifier kernel (P*B = 310 > 256) and leaves the last channel group of the
    weight-gradient kernel partly filled where OC is not a multiple of 4."""
    from biscotti_b200.models.registry import CNN_ARCH, cnn_param_slices
    from biscotti_b200.ops import cnn_native

    arch = CNN_ARCH[kind]
    _, d = cnn_param_slices(arch)
    torch.manual_seed(9)
    P, B = 31, 10
    F = arch["image"][0] * arch["image"][1] * arch["image"][2]
    w = (torch.randn(d) * 0.02).to(DEV)
    Xb = torch.randn(P, B, F, device=DEV)
    yb = torch.randint(0, arch["n_classes"], (P, B), dtype=torch.int32, device=DEV)
    a, b = torch.zeros(P, d + 2, device=DEV), torch.zeros(P, d + 2, device=DEV)
    lib.bsc_cnn_force_generic(1)
    try:
        cnn_native.local_step(kind, w, Xb, yb, 100.0, a)
        torch.cuda.synchronize()
    finally:
        lib.bsc_cnn_force_generic(0)
    cnn_native.local_step(kind, w, Xb, yb, 100.0, b)
    torch.cuda.synchronize()
    scale = float(a.abs().max())
    # (the conv weight gradient of the specialised path is a bf16 hi/lo tensor-core GEMM over 12 960 positions: ~2^-16
    # relative per product, the same bound the autograd comparison uses)
    assert scale > 0 and torch.allclose(a, b, rtol=1e-3, atol=2e-5 * max(1.0, scale)), float((a - b).abs().max())


@pytest.mark.parametrize("kind", ["mnist_cnn", "cifar_cnn"])
def test_cnn_native_eval_matches_torch_forward(kind):
    """Error rate from the native kernels (conv + classifier partial sums + arg-max count) against F.conv2d / F.linear."""
    from biscotti_b200.models.registry import CNN_ARCH, _unflatten, cnn_forward, cnn_param_slices
    from biscotti_b200.ops import cnn_native

    arch = CNN_ARCH[kind]
    _, d = cnn_param_slices(arch)
    torch.manual_seed(10)
    N = 300
    F = arch["image"][0] * arch["image"][1] * arch["image"][2]
    w = (torch.randn(d) * 0.02).to(DEV)
    X = torch.randn(N, F, device=DEV)
    logits = cnn_forward(arch, _unflatten(arch, w), X.view(N, *arch["image"]))
    pred = logits.argmax(-1)
    y = pred.clone().int()
    y[::3] = (y[::3] + 1) % arch["n_classes"]            # a third of the labels made wrong on purpose
    # leave out images whose top two logits are within rounding of each other
    top2 = logits.topk(2, dim=-1).values
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4
    want = float((pred[clear] != y[clear].long()).float().sum())
    got = cnn_native.eval_errors(kind, w, X[clear].contiguous(), y[clear].contiguous())
    assert got is not None
    assert abs(float(got) * int(clear.sum()) - want) < 0.5


@pytest.mark.parametrize("M,N,F,C,with_base", [(1, 2000, 784, 10, False), (70, 2000, 784, 10, True), (9, 333, 100, 2, True),
                                               (17, 128, 64, 16, False)])
def test_eval_on_tensor_cores_matches_fp32_oracle(M, N, F, C, with_base, monkeypatch):
    """eval_tc_kernel (tcgen05 bf16x3 logits GEMM + fused bias / arg-max / error count) against the fp32 PyTorch
    restatement and the SIMT kernel: counts may differ only where two logits tie to within bf16x3 rounding."""
    torch.manual_seed(5)
    d = C * F + C
    protos = torch.randn(C, F)
    y = torch.randint(0, C, (N,), dtype=torch.int32)
    X = protos[y.long()] * 0.5 + torch.randn(N, F)
    base = torch.cat([protos.reshape(-1) * 0.1, torch.zeros(C)]) if with_base else None
    Wc = torch.randn(M, d + 2) * 0.05
    Wc[0, :d] = torch.cat([protos.reshape(-1) * 0.2, torch.randn(C) * 0.1])     # one candidate that actually classifies
    ref = ml.eval_softmax_errors(Wc, base, X, y, C)
    Xg, yg, Wg = X.to(DEV), y.to(DEV), Wc.to(DEV)
    bg = base.to(DEV) if base is not None else None
    monkeypatch.setenv("BISCOTTI_EVAL_IMPL", "tc")
    tc = ml.eval_softmax_errors(Wg, bg, Xg, yg, C).cpu()
    tc2 = ml.eval_softmax_errors(Wg, bg, Xg, yg, C).cpu()                        # cached sample planes, same answer
    monkeypatch.setenv("BISCOTTI_EVAL_IMPL", "simt")
    simt = ml.eval_softmax_errors(Wg, bg, Xg, yg, C).cpu()
    assert torch.equal(tc, tc2)
    assert float((tc - ref).abs().max()) * N <= 2.5, (tc * N, ref * N)
    assert float((simt - ref).abs().max()) * N <= 2.5
    assert float(ref[0]) < 0.5 or C > 10


def test_minibatch_sampler_kernel_draws_distinct_uniform_rows(lib):
    """sample_minibatch_kernel: B distinct rows per peer, a function of (seed, iteration, peer) only, uniform."""
    from biscotti_b200.ops._call import ptr, stream

    P, Ns, B = 64, 600, 10
    ids = torch.arange(100, 100 + P, dtype=torch.int32, device=DEV)
    out = torch.empty(P, B, dtype=torch.int32, device=DEV)

    def draw(seed, it, pid=ids, n=P):
        assert lib.bsc_sample_minibatch(seed, it, ptr(pid), n, Ns, B, ptr(out), stream(torch.device(DEV))) == 0
        return out[:n].cpu().clone()

    a = draw(7, 3)
    assert int(a.min()) >= 0 and int(a.max()) < Ns
    assert all(len(set(r.tolist())) == B for r in a)
    assert torch.equal(a, draw(7, 3)) and not torch.equal(a, draw(7, 4)) and not torch.equal(a, draw(8, 3))
    sub = draw(7, 3, ids[5:9].contiguous(), 4)                      # a peer's rows do not depend on who else is drawn
    assert torch.equal(sub, a[5:9])
    counts = torch.zeros(Ns)
    for it in range(200):
        counts += torch.bincount(draw(1, it).flatten().long(), minlength=Ns).float()
    expect = 200 * P * B / Ns
    assert float((counts - expect).abs().max()) < 6 * expect ** 0.5   # every row is drawn about equally often
    assert lib.bsc_sample_minibatch(1, 0, ptr(ids), P, 2000, B, ptr(out), stream(torch.device(DEV))) == -1
