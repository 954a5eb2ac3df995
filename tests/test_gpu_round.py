"""Persistent FL-round kernel (cs_tiny_ecg_train_round): one cooperative launch = all local steps of a round
(+ the FedAvg aggregation).  Oracles: the single-launch step kernel, fp32 PyTorch autograd + torch SGD, and the
fp64 weighted mean for the fused aggregation (SURVEY 6.3 b/c).  Needs a B200."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


def _need_sm100():
    from crossscale.ops import _lib
    if not _lib.is_sm100():
        pytest.skip("needs sm_100")


def _model(Cn=2, seed=0):
    from crossscale.models import TinyECG
    torch.manual_seed(seed)
    m = TinyECG(Cn).to(_dev())
    with torch.no_grad():
        for p in m.parameters():
            p.mul_(2.0)
    return m


def _data(n, L, Cn, seed=1, dtype=torch.float32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, L, device=_dev(), generator=g).to(dtype)
    y = torch.randint(0, Cn, (n,), device=_dev(), generator=g)
    return x, y


def _rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


def _state(m, Cn, amp_name, **kw):
    from crossscale.ops import tiny_ecg as tk
    flat = m.flatten_parameters().flat.clone()
    return tk.TinyTrainState(flat, Cn, lr=1e-2, momentum=0.9, amp=tk.AMP_BY_NAME[amp_name],
                             scaler=tk.ScalerConfig(enabled=(amp_name == "fp16"), **kw))


@pytest.mark.parametrize("amp_name", ["bf16", "fp16"])
@pytest.mark.parametrize("Cn,B,L,steps", [(2, 256, 500, 12), (5, 37, 360, 7), (2, 400, 500, 4), (3, 9, 1000, 5), (2, 64, 129, 1)])
def test_round_matches_single_launch_steps(amp_name, Cn, B, L, steps):
    """Same batches through ONE persistent launch and through `steps` launches of the step kernel: the two differ
    only in the order fp32 partial sums meet (atomics), so parameters agree to ~1e-5 relative."""
    _need_sm100()
    from crossscale.ops import tiny_ecg as tk
    m = _model(Cn, seed=5)
    x, y = _data(3000, L, Cn)
    idx = torch.randint(0, x.shape[0], (steps * B,), device=_dev())
    a, b = _state(m, Cn, amp_name), _state(m, Cn, amp_name)
    for s in range(steps):
        a.step(x, y, idx[s * B:(s + 1) * B], batch=B)
    b.round(x, y, steps, B, idx=idx, sample_mode=tk.SAMPLE_IDX)
    torch.cuda.synchronize()
    b.check_errors()
    n = 1424 + 17 * Cn
    assert _rel(b.params[:n], a.params[:n]) < 2e-5, _rel(b.params[:n], a.params[:n])
    assert _rel(b.mom[:n], a.mom[:n]) < 2e-4, _rel(b.mom[:n], a.mom[:n])
    assert abs(b.loss_sum() - a.loss_sum()) < 1e-4 * max(1.0, abs(a.loss_sum()))
    assert b.steps_done() == steps and a.steps_done() == steps
    assert float(b.fstate[tk.FS_LAST_LOSS]) == pytest.approx(float(a.fstate[tk.FS_LAST_LOSS]), rel=1e-3, abs=1e-5)
    assert b.loss_scale() == a.loss_scale()
    assert int(b.ctrl[tk.CTRL_TICKET]) == 0
    assert float(b.gacc_round.abs().max()) == 0.0          # the rotating scratch is left clean
    # a second round on the same state keeps working (barrier counter and scratch were reset)
    b.round(x, y, steps, B, idx=idx, sample_mode=tk.SAMPLE_IDX)
    for s in range(steps):
        a.step(x, y, idx[s * B:(s + 1) * B], batch=B)
    torch.cuda.synchronize()
    assert _rel(b.params[:n], a.params[:n]) < 5e-5


def test_round_tracks_fp32_torch_sgd():
    """fp32 oracle: stock nn.Module + F.cross_entropy + torch.optim.SGD(1e-2, 0.9) on the same batches."""
    _need_sm100()
    from crossscale.models import TinyECG
    from crossscale.ops import tiny_ecg as tk
    Cn, B, L, steps = 2, 128, 500, 6
    m = _model(Cn, seed=3)
    ref = TinyECG(Cn).to(_dev())
    ref.load_state_dict(m.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr=1e-2, momentum=0.9)
    x, y = _data(2000, L, Cn, seed=9)
    idx = torch.randint(0, x.shape[0], (steps * B,), device=_dev())
    st = _state(m, Cn, "bf16")
    st.round(x, y, steps, B, idx=idx)
    losses = []
    for s in range(steps):
        sel = idx[s * B:(s + 1) * B]
        opt.zero_grad(set_to_none=True)
        loss = F.cross_entropy(ref.forward_reference(x[sel].unsqueeze(1)), y[sel])
        loss.backward()
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    st.check_errors()
    want = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    err = _rel(st.params[:want.numel()], want)
    assert err < 5e-3, err                                  # bf16 operands, 6 steps
    assert st.loss_sum() == pytest.approx(sum(losses), rel=2e-2)


def test_round_in_kernel_permutation_sampler():
    """sample_mode 2 draws the rows the host mirror predicts, across an epoch boundary, and advances its counter."""
    _need_sm100()
    from crossscale.ops import tiny_ecg as tk
    Cn, B, L, steps, n = 2, 64, 500, 5, 300                 # usable = 256 rows/epoch: 5 steps of 64 cross into epoch 1
    m = _model(Cn, seed=8)
    x, y = _data(n, L, Cn, seed=4)
    a, b = _state(m, Cn, "bf16"), _state(m, Cn, "bf16")
    seed = 1234
    rows = tk.perm_batches(0, steps * B, n, B, seed)
    assert sorted(rows[:256]) == sorted(set(rows[:256])) and len(set(rows[:256])) == 256   # an epoch never repeats a row
    idx = torch.tensor(rows, dtype=torch.int64, device=_dev())
    a.round(x, y, steps, B, idx=idx, sample_mode=tk.SAMPLE_IDX)
    b.round(x, y, steps, B, sample_mode=tk.SAMPLE_PERM, seed=seed)
    torch.cuda.synchronize()
    n_par = 1424 + 17 * Cn
    assert _rel(b.params[:n_par], a.params[:n_par]) < 2e-5
    assert b.samples_drawn() == steps * B
    rows2 = tk.perm_batches(steps * B, steps * B, n, B, seed)
    a.round(x, y, steps, B, idx=torch.tensor(rows2, dtype=torch.int64, device=_dev()), sample_mode=tk.SAMPLE_IDX)
    b.round(x, y, steps, B, sample_mode=tk.SAMPLE_PERM, seed=seed)
    torch.cuda.synchronize()
    assert _rel(b.params[:n_par], a.params[:n_par]) < 5e-5
    assert b.samples_drawn() == 2 * steps * B


def test_round_staged_rows_bf16_input():
    _need_sm100()
    from crossscale.ops import tiny_ecg as tk
    Cn, B, L, steps = 2, 96, 500, 4
    m = _model(Cn, seed=2)
    x, y = _data(steps * B, L, Cn, seed=6, dtype=torch.bfloat16)
    a, b = _state(m, Cn, "bf16"), _state(m, Cn, "bf16")
    for s in range(steps):
        a.step(x[s * B:(s + 1) * B], y[s * B:(s + 1) * B], None, batch=B)
    b.round(x, y, steps, B, sample_mode=tk.SAMPLE_ROWS)
    torch.cuda.synchronize()
    n = 1424 + 17 * Cn
    assert _rel(b.params[:n], a.params[:n]) < 2e-5


def test_round_scaler_skips_and_backs_off_like_the_step_kernel():
    """fp16 + GradScaler inside the persistent kernel: an absurd initial scale overflows, steps are skipped and the
    scale halves until gradients are finite - step for step like the single-launch kernel."""
    _need_sm100()
    from crossscale.ops import tiny_ecg as tk
    Cn, B, L, steps = 2, 64, 500, 10
    m = _model(Cn, seed=7)
    x, y = _data(1000, L, Cn, seed=3)
    idx = torch.randint(0, x.shape[0], (steps * B,), device=_dev())
    kw = dict(init_scale=2.0 ** 30, growth_interval=3)
    a, b = _state(m, Cn, "fp16", **kw), _state(m, Cn, "fp16", **kw)
    for s in range(steps):
        a.step(x, y, idx[s * B:(s + 1) * B], batch=B)
    b.round(x, y, steps, B, idx=idx)
    torch.cuda.synchronize()
    assert a.skipped_steps() > 0
    assert b.skipped_steps() == a.skipped_steps()
    assert b.loss_scale() == a.loss_scale()
    assert int(b.ctrl[tk.CTRL_GROWTH]) == int(a.ctrl[tk.CTRL_GROWTH])
    n = 1424 + 17 * Cn
    assert _rel(b.params[:n], a.params[:n]) < 5e-5


def test_out_of_range_label_raises():
    _need_sm100()
    from crossscale.ops import _lib, tiny_ecg as tk
    Cn, B, L = 2, 32, 500
    m = _model(Cn)
    x, y = _data(B, L, Cn)
    y[5] = 4                                                # a 5-class sidecar fed to a 2-class model
    with pytest.raises(ValueError):
        tk.validate_labels(y, Cn)
    for amp_name in ("bf16", "none"):
        st = _state(m, Cn, amp_name)
        st.step(x, y, None, batch=B)
        torch.cuda.synchronize()
        with pytest.raises(_lib.NativeError):
            st.check_errors()
    st = _state(m, Cn, "bf16")
    st.round(x, y, 2, B, sample_mode=tk.SAMPLE_IDX,
             idx=torch.arange(B, device=_dev()).repeat(2))
    torch.cuda.synchronize()
    with pytest.raises(_lib.NativeError):
        st.check_errors()
    assert torch.isfinite(st.params).all()


@pytest.mark.parametrize("world,wire", [(2, "fp32"), (4, "fp32"), (4, "bf16")])
def test_round_fused_aggregation_emulated_clients(world, wire):
    """`world` clients on one GPU, ONE cooperative launch: local steps + in-kernel one-shot FedAvg push.
    Oracle: each client trained alone with the step kernel, then sum_i w_i theta_i in fp64."""
    _need_sm100()
    from crossscale.ops import _lib, tiny_ecg as tk
    from crossscale.parallel.symm import LocalGroup, ptr_array
    Cn, B, L, steps = 2, 32, 500, 3
    lib = _lib.cuda_lib()
    m = _model(Cn, seed=12)
    n = 1424 + 17 * Cn
    counts = [1000 + 300 * r for r in range(world)]
    wts = [c / sum(counts) for c in counts]
    xs, ys, idxs, solo, fused, prevs = [], [], [], [], [], []
    for r in range(world):
        x, y = _data(600, L, Cn, seed=20 + r)
        xs.append(x); ys.append(y)
        idxs.append(torch.randint(0, 600, (steps * B,), device=_dev()))
        solo.append(_state(m, Cn, "bf16")); fused.append(_state(m, Cn, "bf16"))
        prevs.append(fused[r].params.clone())
    group = LocalGroup(world, int(lib.cs_fedavg_arena_bytes(n, int(wire == "bf16"), world)))
    for rounds in range(2):                                 # two rounds: epoch / parity reuse of the arena slots
        start = [p.clone() for p in prevs]
        for r in range(world):
            solo[r].params.copy_(start[r])
            for s in range(steps):
                solo[r].step(xs[r], ys[r], idxs[r][s * B:(s + 1) * B], batch=B)
        for st in fused:
            if st.gacc_round is None:
                st.gacc_round = torch.zeros(int(lib.cs_tiny_ecg_round_scratch_floats(Cn)), device=_dev())
        w = (C.c_float * world)(*wts)
        nrows = (C.c_int64 * world)(*[600] * world)
        seeds = (C.c_uint64 * world)(*[0] * world)
        _lib.check(lib.cs_tiny_ecg_train_round_local(
            group.handles, world, ptr_array(xs), ptr_array(ys), ptr_array(idxs), 0, nrows, B, L, Cn, steps, tk.SAMPLE_IDX,
            seeds, ptr_array([s.params for s in fused]), ptr_array([s.mom for s in fused]),
            ptr_array([s.gacc_round for s in fused]), ptr_array([s.ctrl for s in fused]),
            ptr_array([s.fstate for s in fused]), 1e-2, 0.9, 0.0, tk.AMP_BF16, 0, 2.0, 0.5, 2000, ptr_array(prevs), w,
            int(wire == "bf16"), 1.0, 0, 0, 0, _lib.stream_ptr()), "cs_tiny_ecg_train_round_local")
        torch.cuda.synchronize()
        assert group.status() == 0
        for st in fused:
            st.check_errors()
        want = start[0][:n].double() + sum(wts[r] * (solo[r].params[:n].double() - start[r][:n].double()) for r in range(world))
        tol = 3e-5 if wire == "fp32" else 2e-3
        for r in range(world):
            assert _rel(fused[r].params[:n], want) < tol, (r, _rel(fused[r].params[:n], want))
            assert torch.equal(fused[r].params[:n], fused[0].params[:n])      # bit-identical replicas
            assert torch.equal(prevs[r][:n], fused[r].params[:n])
            assert _rel(fused[r].mom[:n], solo[r].mom[:n]) < 5e-4             # momentum stays local
            solo[r].mom.copy_(fused[r].mom)
    group.close()


# ---- G0: fp32 operands on the tensor core as TF32 (tiny_ecg_round_tf32.cu) -----------------------------------------------------
@pytest.mark.parametrize("Cn,B,L,steps", [(2, 256, 500, 8), (5, 37, 360, 7), (2, 400, 500, 3), (3, 9, 129, 4), (2, 64, 512, 1), (8, 3, 64, 2)])
def test_round_tf32_matches_fp32_step_kernel(Cn, B, L, steps):
    """amp = fp32 through the persistent round (tcgen05 kind::tf32, fp32 accumulation, conv1/head/SGD in fp32 FFMA) against
    `steps` launches of the fp32 FFMA step kernel on the same batches.  TF32 keeps 10 mantissa bits of each conv2 operand, so
    the two agree to ~1e-3 of the update, far tighter than bf16 (8 bits)."""
    _need_sm100()
    from crossscale.ops import tiny_ecg as tk
    m = _model(Cn, seed=5)
    x, y = _data(3000, L, Cn)
    idx = torch.randint(0, x.shape[0], (steps * B,), device=_dev())
    a, b = _state(m, Cn, "none"), _state(m, Cn, "none")
    start = a.params.clone()
    for s in range(steps):
        a.step(x, y, idx[s * B:(s + 1) * B], batch=B)
    assert b.round_supported(L)
    b.round(x, y, steps, B, idx=idx, sample_mode=tk.SAMPLE_IDX)
    torch.cuda.synchronize()
    b.check_errors()
    n = 1424 + 17 * Cn
    upd = _rel(b.params[:n] - start[:n], a.params[:n] - start[:n])
    assert upd < 4e-3, upd                                   # error relative to the UPDATE, not to the parameters
    assert _rel(b.params[:n], a.params[:n]) < 2e-4
    assert abs(b.loss_sum() - a.loss_sum()) < 1e-3 * max(1.0, abs(a.loss_sum()))
    assert b.steps_done() == steps
    assert int(b.ctrl[tk.CTRL_TICKET]) == 0 and float(b.gacc_round.abs().max()) == 0.0
    b.round(x, y, steps, B, idx=idx, sample_mode=tk.SAMPLE_IDX)              # a second round on the same state
    for s in range(steps):
        a.step(x, y, idx[s * B:(s + 1) * B], batch=B)
    torch.cuda.synchronize()
    b.check_errors()
    assert _rel(b.params[:n], a.params[:n]) < 4e-4


def test_round_tf32_tracks_fp32_torch_sgd_and_beats_bf16():
    """fp32 oracle (nn.Module + F.cross_entropy + torch.optim.SGD, TF32 off): the tf32 round must sit closer to it than the
    bf16 round on the same batches."""
    _need_sm100()
    from crossscale.models import TinyECG
    Cn, B, L, steps = 5, 128, 360, 6
    m = _model(Cn, seed=3)
    ref = TinyECG(Cn).to(_dev())
    ref.load_state_dict(m.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr=1e-2, momentum=0.9)
    x, y = _data(2000, L, Cn, seed=9)
    idx = torch.randint(0, x.shape[0], (steps * B,), device=_dev())
    st32, st16 = _state(m, Cn, "none"), _state(m, Cn, "bf16")
    st32.round(x, y, steps, B, idx=idx)
    st16.round(x, y, steps, B, idx=idx)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        losses = []
        for s in range(steps):
            sel = idx[s * B:(s + 1) * B]
            opt.zero_grad(set_to_none=True)
            loss = F.cross_entropy(ref.forward_reference(x[sel].unsqueeze(1)), y[sel])
            loss.backward()
            opt.step()
            losses.append(float(loss))
    finally:
        torch.backends.cudnn.allow_tf32 = old
    torch.cuda.synchronize()
    st32.check_errors()
    want = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    e32, e16 = _rel(st32.params[:want.numel()], want), _rel(st16.params[:want.numel()], want)
    assert e32 < 5e-4, e32
    assert e32 < e16, (e32, e16)
    assert st32.loss_sum() == pytest.approx(sum(losses), rel=2e-3)


def test_round_tf32_permutation_sampler_and_fused_aggregation():
    """the tf32 kernel shares the sampler and the FedAvg tail with the bf16 one: 3 emulated clients, in-kernel permutation."""
    _need_sm100()
    from crossscale.ops import _lib, tiny_ecg as tk
    from crossscale.parallel.symm import LocalGroup, ptr_array
    Cn, B, L, steps, world, rows = 2, 48, 360, 3, 3, 500
    lib = _lib.cuda_lib()
    m = _model(Cn, seed=12)
    n = 1424 + 17 * Cn
    wts = [0.5, 0.3, 0.2]
    xs, ys, solo, fused, prevs = [], [], [], [], []
    for r in range(world):
        x, y = _data(rows, L, Cn, seed=20 + r)
        xs.append(x); ys.append(y)
        solo.append(_state(m, Cn, "none")); fused.append(_state(m, Cn, "none"))
        prevs.append(fused[r].params.clone())
        fused[r].gacc_round = torch.zeros(int(lib.cs_tiny_ecg_round_scratch_floats(Cn)), device=_dev())
    start = prevs[0].clone()
    for r in range(world):
        idx = torch.tensor(tk.perm_batches(0, steps * B, rows, B, 77 + r), dtype=torch.int64, device=_dev())
        for s in range(steps):
            solo[r].step(xs[r], ys[r], idx[s * B:(s + 1) * B], batch=B)
    group = LocalGroup(world, int(lib.cs_fedavg_arena_bytes(n, 0, world)))
    w = (C.c_float * world)(*wts)
    nrows = (C.c_int64 * world)(*[rows] * world)
    seeds = (C.c_uint64 * world)(*[77 + r for r in range(world)])
    _lib.check(lib.cs_tiny_ecg_train_round_local(
        group.handles, world, ptr_array(xs), ptr_array(ys), None, 0, nrows, B, L, Cn, steps, tk.SAMPLE_PERM,
        seeds, ptr_array([s.params for s in fused]), ptr_array([s.mom for s in fused]),
        ptr_array([s.gacc_round for s in fused]), ptr_array([s.ctrl for s in fused]),
        ptr_array([s.fstate for s in fused]), 1e-2, 0.9, 0.0, tk.AMP_FP32, 0, 2.0, 0.5, 2000, ptr_array(prevs), w,
        0, 1.0, 0, 0, 0, _lib.stream_ptr()), "cs_tiny_ecg_train_round_local")
    torch.cuda.synchronize()
    assert group.status() == 0
    for st in fused:
        st.check_errors()
    want = start[:n].double() + sum(wts[r] * (solo[r].params[:n].double() - start[:n].double()) for r in range(world))
    for r in range(world):
        assert _rel(fused[r].params[:n], want) < 2e-4, (r, _rel(fused[r].params[:n], want))
        assert torch.equal(fused[r].params[:n], fused[0].params[:n])
        assert fused[r].samples_drawn() == steps * B
    group.close()
