"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the reference's
golden vectors.  Run on the B200 box: ``pytest -m gpu``.

Tolerances (BASELINE.json north_star): probabilities within 2e-3 and argmax identical on
>= 99.9 % of queries for the 16-bit tensor-core path; within 1e-5 for the fp32 path; the
ensemble mean equals float32(reference) exactly.

The tensor-core path runs on fp16 operands by default (RFK_F16).  With bf16 operands
(RFK_BF16) the same kernel measures 3e-3..7e-3 and 99.5..99.9 % on these networks -- bf16's
8-bit significand alone exceeds the 2e-3 gate, see DESIGN.md "Operand format" -- so bf16 is
tested against a documented looser bound.
"""
import os

import numpy as np
import pytest
import torch

from oracle import feedforward as ff
from rafiki_b200 import _native, engine, synth

pytestmark = pytest.mark.gpu

TOL_TC = 2e-3          # fp16 operands, fp32 accumulate
TOL_BF16 = 1.5e-2      # bf16 operands (optional mode; see module docstring)
TOL_F32 = 1e-5


@pytest.fixture(scope='module', autouse=True)
def _device():
    assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
    assert _native.load().rfk_device_check(0) == 0, _native.load().rfk_last_error()
    torch.cuda.set_device(0)
    yield
    torch.cuda.synchronize()


def _np(t):
    return t.detach().float().cpu().numpy()


# ------------------------------------------------------------------------------- kernel 1
@pytest.mark.parametrize('n,d', [(1, 784), (333, 784), (4096, 784), (64, 3072), (7, 1024)])
@pytest.mark.parametrize('out_dtype', [torch.float32, torch.bfloat16])
def test_decode_affine_vectorised(n, d, out_dtype):
    rng = np.random.default_rng(n + d)
    X = rng.integers(0, 256, size=(n, d), dtype=np.uint8)
    p = synth.make_trial(1, int(round((d if d != 3072 else 1024) ** 0.5)), 3 if d == 3072 else 1, 1, 8,
                         dead=synth.dead_mask(d))
    inv, shift = ff.bn_affine(p)
    want = X.astype(np.float32) * inv + shift
    got = engine.decode_u8(torch.from_numpy(X).cuda(), torch.from_numpy(inv).cuda(), torch.from_numpy(shift).cuda(),
                           out_dtype)
    if out_dtype == torch.float32:
        # fused multiply-add vs NumPy's multiply-then-add: one rounding apart
        np.testing.assert_allclose(_np(got), want, rtol=3e-7, atol=1e-6)
    else:
        ref = torch.from_numpy(want).to(torch.bfloat16).float().numpy()
        np.testing.assert_allclose(_np(got), ref, rtol=2 ** -7, atol=1e-6)
        assert np.mean(_np(got) == ref) > 0.999
    # plain cast is exact in both dtypes
    got = engine.decode_u8(torch.from_numpy(X).cuda(), None, None, out_dtype)
    assert np.array_equal(_np(got), X.astype(np.float32))


def test_decode_ragged_and_padded():
    rng = np.random.default_rng(5)
    X = rng.integers(0, 256, size=(37, 75), dtype=np.uint8)     # d not a multiple of 16
    sc = rng.normal(size=75).astype(np.float32)
    sh = rng.normal(size=75).astype(np.float32)
    got = engine.decode_u8(torch.from_numpy(X).cuda(), torch.from_numpy(sc).cuda(), torch.from_numpy(sh).cuda(),
                           torch.float32, ld_out=80)
    want = np.zeros((37, 80), np.float32)
    want[:, :75] = X.astype(np.float32) * sc + sh
    np.testing.assert_allclose(_np(got), want, rtol=3e-7, atol=1e-6)
    assert engine.decode_u8(torch.empty((0, 784), dtype=torch.uint8, device='cuda')).shape == (0, 784)


def test_resize_nearest_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, 'resize_reference.npz'))
    x = torch.from_numpy(g['queries']).cuda()
    got = engine.resize_decode_u8(x, 32)
    assert np.array_equal(got.cpu().numpy().reshape(-1, 32, 32), g['nearest_32'])
    assert np.array_equal(engine.resize_decode_u8(x, 28).cpu().numpy().reshape(-1, 28, 28), g['identity_28'])
    rgb = torch.from_numpy(g['rgb_in']).cuda()
    assert np.array_equal(engine.resize_decode_u8(rgb, 32).cpu().numpy().reshape(-1, 32, 32, 3), g['rgb_identity_32'])
    f = engine.resize_decode_u8(x, 32, out_dtype=torch.float32)
    assert np.array_equal(_np(f).reshape(-1, 32, 32), g['nearest_32'].astype(np.float32))


# ------------------------------------------------------------------------------- kernel 3
@pytest.mark.parametrize('name', ['w1', 'w2', 'w3', 'w8', 'w5c3', 'onehot'])
def test_ensemble_mean_equals_reference_rounded_once(golden_dir, name):
    g = np.load(os.path.join(golden_dir, 'ensemble_reference.npz'))
    p, want = g[name + '_in'], g[name + '_out']
    got = engine.ensemble_mean([torch.from_numpy(np.ascontiguousarray(q)).cuda() for q in p])
    assert np.array_equal(got.cpu().numpy(), want.astype(np.float32))
    assert np.abs(got.cpu().numpy().astype(np.float64) - want).max() <= 1e-6


def test_ensemble_mean_large_and_list_api():
    rng = np.random.default_rng(3)
    p = rng.random((5, 70001, 10), dtype=np.float32)
    got = engine.ensemble_mean([torch.from_numpy(q).cuda() for q in p])
    assert np.array_equal(got.cpu().numpy(), ff.ensemble_mean_array(p).astype(np.float32))
    from rafiki_b200.predictor.ensemble import ensemble_predictions
    small = p[:3, :9].tolist()
    out = ensemble_predictions(small, 'IMAGE_CLASSIFICATION')
    ref = ff.ensemble_predictions(small, 'IMAGE_CLASSIFICATION')
    assert isinstance(out, list) and isinstance(out[0][0], float)
    assert np.array_equal(np.asarray(out), np.asarray(ref).astype(np.float32).astype(np.float64))


# ------------------------------------------------------------------------------- kernel 2
SHAPES = [(28, 1, 2, 128), (28, 1, 1, 36), (28, 1, 2, 7), (32, 1, 2, 36), (28, 1, 1, 2), (28, 1, 2, 100),
          (28, 1, 1, 128), (32, 3, 2, 64)]


def _case(S, ch, L, H, n, t=0, kind='fashion'):
    d = S * S * ch
    dead = synth.dead_mask(d)
    params = synth.make_trial(t, S, ch, L, H, dead=dead, data_kind=kind)
    X = synth.make_queries(n, S, ch, kind, seed=t, dead=dead)
    return params, X


@pytest.mark.parametrize('S,ch,L,H', SHAPES)
def test_trial_forward_fp32_path(S, ch, L, H):
    params, X = _case(S, ch, L, H, 517)
    tr = engine.Trial(params, compute='f32')
    got = _np(tr.forward(torch.from_numpy(X).cuda()))
    want = ff.forward(params, X, np.float32)
    assert np.abs(got - want).max() <= TOL_F32
    assert np.abs(got - ff.forward(params, X, np.float64)).max() <= TOL_F32
    tr.close()


@pytest.mark.parametrize('S,ch,L,H', SHAPES)
@pytest.mark.parametrize('n', [1, 128, 517, 4096])
def test_trial_forward_tensor_core_path(S, ch, L, H, n):
    params, X = _case(S, ch, L, H, n)
    tr = engine.Trial(params, compute='f16')
    assert tr.fused
    got = _np(tr.forward(torch.from_numpy(X).cuda()))
    want = ff.forward(params, X, np.float32)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= TOL_TC
    assert np.abs(got.sum(1) - 1).max() <= 1e-5
    tr.close()


def test_argmax_agreement_over_all_shapes():
    """>= 99.9 % identical argmax, pooled over the shape list at 4096 queries each (a flip
    needs an oracle top-2 gap below twice the probability tolerance)."""
    same = total = 0
    for (S, ch, L, H) in SHAPES:
        params, X = _case(S, ch, L, H, 4096, t=5)
        tr = engine.Trial(params, compute='f16')
        got = _np(tr.forward(torch.from_numpy(X).cuda()))
        want = ff.forward(params, X, np.float32)
        flips = got.argmax(1) != want.argmax(1)
        top2 = np.sort(want, axis=1)[:, -2:]
        assert np.all((top2[flips, 1] - top2[flips, 0]) <= 2 * TOL_TC)
        same += int((~flips).sum())
        total += len(flips)
        tr.close()
    assert same / total >= 0.999


@pytest.mark.parametrize('S,ch,L,H', [(28, 1, 2, 128), (28, 1, 1, 36), (32, 3, 2, 64)])
def test_trial_forward_bf16_mode(S, ch, L, H):
    params, X = _case(S, ch, L, H, 1000)
    tr = engine.Trial(params, compute='bf16')
    assert tr.fused
    got = _np(tr.forward(torch.from_numpy(X).cuda()))
    want = ff.forward(params, X, np.float32)
    assert np.abs(got - want).max() <= TOL_BF16
    assert np.mean(got.argmax(1) == want.argmax(1)) >= 0.99
    tr.close()


def test_fp16_weight_range_is_checked():
    params, X = _case(28, 1, 1, 8, 4)
    params['kernels'][1] = params['kernels'][1] * 1e6
    with pytest.raises(_native.RafikiNativeError, match='does not fit fp16'):
        engine.Trial(params, compute='f16')
    engine.Trial(params, compute='bf16').close()


def test_sample_query_of_the_reference(golden_dir):
    q = np.load(os.path.join(golden_dir, 'resize_reference.npz'))['sample_query']
    g = np.load(os.path.join(golden_dir, 'forward_restatement.npz'))
    dead = synth.dead_mask(784)
    for t, (L, H) in enumerate([(2, 128), (1, 36), (2, 7)]):
        params = synth.make_trial(t, 28, 1, L, H, dead=dead, data_kind='fashion')
        X = synth.make_queries(16, 28, 1, 'fashion', seed=t, dead=dead)
        X[0] = q
        want = g['unpinned_t{}_probs_f64'.format(t)]
        for compute, tol in (('f16', TOL_TC), ('bf16', TOL_BF16), ('f32', TOL_F32)):
            tr = engine.Trial(params, compute=compute)
            assert np.abs(_np(tr.forward(torch.from_numpy(X).cuda())) - want).max() <= tol
            tr.close()


def test_zero_weights_give_uniform_probabilities():
    params, X = _case(28, 1, 2, 16, 64)
    params['kernels'] = [np.zeros_like(k) for k in params['kernels']]
    params['biases'] = [np.zeros_like(b) for b in params['biases']]
    for compute in ('f16', 'bf16', 'f32'):
        tr = engine.Trial(params, compute=compute)
        assert np.allclose(_np(tr.forward(torch.from_numpy(X).cuda())), 0.1, atol=1e-6)
        tr.close()


# ------------------------------------------------------------------------------- kernels 1+2+3
@pytest.mark.parametrize('compute,tol', [('f16', TOL_TC), ('bf16', TOL_BF16), ('f32', TOL_F32)])
@pytest.mark.parametrize('w,n', [(3, 4096), (8, 1000), (2, 129), (1, 64)])
def test_ensemble_forward(compute, tol, w, n):
    dead = synth.dead_mask(784)
    shapes = [(2, 128), (1, 36), (2, 7), (2, 128), (1, 128), (2, 64), (1, 2), (2, 100)]
    params = [synth.make_trial(t, 28, 1, *shapes[t], dead=dead, data_kind='fashion') for t in range(w)]
    X = synth.make_queries(n, 28, 1, 'fashion', seed=11, dead=dead)
    x = torch.from_numpy(X).cuda()
    trials = [engine.Trial(p, compute=compute) for p in params]
    ens = engine.Ensemble(trials)
    got = ens.forward(x).cpu().numpy()
    want = ff.ensemble_mean_array([ff.forward(p, X, np.float32) for p in params])
    assert np.abs(got.astype(np.float64) - want).max() <= tol
    # the fused kernel equals "each trial alone, then kernel 3" bit for bit
    singles = [t.forward(x) for t in trials]
    assert np.array_equal(got, engine.ensemble_mean(singles).cpu().numpy())
    if n >= 4096 and compute != 'bf16':
        assert np.mean(got.argmax(1) == want.argmax(1)) >= 0.999
    # a rank holding part of a larger ensemble: partial sums add up to the full mean
    if w >= 2:
        a = engine.Ensemble(trials[:1], total_trials=w).forward(x).cpu().numpy().astype(np.float64)
        b = engine.Ensemble(trials[1:], total_trials=w).forward(x).cpu().numpy().astype(np.float64)
        assert np.abs((a + b) - got).max() <= 1e-6
    for t in trials:
        t.close()


@pytest.mark.parametrize('w,n', [(3, 4096), (8, 1000), (2, 129)])
def test_ensemble_forward_whole_tile_schedule(w, n, monkeypatch):
    """The same small cases with the one-item-per-SM ("split") schedule switched off, so that they run
    the schedule large batches run: a CTA takes whole tiles, all trials of a tile back to back, the
    ensemble sum in registers.  Both schedules must give the same bits."""
    dead = synth.dead_mask(784)
    shapes = [(2, 128), (1, 36), (2, 7), (2, 128), (1, 128), (2, 64), (1, 2), (2, 100)]
    params = [synth.make_trial(t, 28, 1, *shapes[t], dead=dead, data_kind='fashion') for t in range(w)]
    X = synth.make_queries(n, 28, 1, 'fashion', seed=11, dead=dead)
    x = torch.from_numpy(X).cuda()
    trials = [engine.Trial(p) for p in params]
    ens = engine.Ensemble(trials)
    split = ens.forward(x).cpu().numpy()
    monkeypatch.setenv('RFK_NO_SPLIT', '1')
    whole = ens.forward(x).cpu().numpy()
    monkeypatch.delenv('RFK_NO_SPLIT')
    want = ff.ensemble_mean_array([ff.forward(p, X, np.float32) for p in params])
    assert np.abs(whole.astype(np.float64) - want).max() <= TOL_TC
    assert np.array_equal(whole, split)
    for t in trials:
        t.close()


def test_whole_tile_schedule_heterogeneous_trials_batch_20000():
    """Mid-size batch (157 tiles x 3 trials > 148 SMs: whole-tile schedule, several tiles per CTA)
    with trials of different depth and width, against the oracle on every row."""
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, L, H, dead=dead, data_kind='fashion')
              for t, (L, H) in enumerate([(2, 128), (1, 36), (2, 77)])]
    X = synth.make_queries(20000, 28, 1, 'fashion', seed=21, dead=dead)
    x = torch.from_numpy(X).cuda()
    trials = [engine.Trial(p) for p in params]
    got = engine.Ensemble(trials).forward(x).cpu().numpy()
    want = ff.ensemble_mean_array([ff.forward(p, X, np.float32) for p in params])
    assert np.abs(got.astype(np.float64) - want).max() <= TOL_TC
    assert np.mean(got.argmax(1) == want.argmax(1)) >= 0.999
    # equals "each trial alone, then kernel 3" bit for bit
    singles = [t.forward(x) for t in trials]
    assert np.array_equal(got, engine.ensemble_mean(singles).cpu().numpy())
    for t in trials:
        t.close()


@pytest.mark.parametrize('S,n,w', [(28, 20000, 3), (32, 9700, 2), (28, 9473, 1)])
def test_pair_kernel_equals_the_single_cta_kernel(S, n, w, monkeypatch):
    """Large batches run on CTA pairs (tcgen05 cta_group::2, weights resident in shared memory, trial-outer
    order, kernel 3 over per-trial outputs); RFK_NO_PAIR=1 sends the same call to the single-CTA kernel
    (weights streamed, ensemble sum in registers).  Same arithmetic, same bits -- and both against the oracle."""
    d = S * S
    dead = synth.dead_mask(d)
    shapes = [(2, 128), (1, 36), (2, 77)]
    params = [synth.make_trial(t, S, 1, *shapes[t], dead=dead, data_kind='fashion') for t in range(w)]
    X = synth.make_queries(n, S, 1, 'fashion', seed=31, dead=dead)
    x = torch.from_numpy(X).cuda()
    trials = [engine.Trial(p) for p in params]
    ens = engine.Ensemble(trials)
    before = _native.launch_count('fused_pair')
    pair = ens.forward(x).clone()
    assert _native.launch_count('fused_pair') == before + 1          # the pair kernel really ran
    monkeypatch.setenv('RFK_NO_PAIR', '1')
    single = ens.forward(x).clone()
    monkeypatch.delenv('RFK_NO_PAIR')
    assert _native.launch_count('fused_pair') == before + 1
    assert torch.equal(pair, single)
    sl = slice(n - 600, n)
    want = ff.ensemble_mean_array([ff.forward(p, X[sl], np.float32) for p in params])
    assert np.abs(pair[sl].cpu().numpy().astype(np.float64) - want).max() <= TOL_TC
    # one trial of a larger ensemble (what a rank holds when trials are placed one per GPU): float(p / total),
    # the same bits from both kernels for a power-of-two total and for an odd one
    for total in (8, 3):
        part = engine.Ensemble(trials[:1], total_trials=total)
        a = part.forward(x).clone()
        monkeypatch.setenv('RFK_NO_PAIR', '1')
        b = part.forward(x).clone()
        monkeypatch.delenv('RFK_NO_PAIR')
        assert torch.equal(a, b), 'total_trials = {}'.format(total)
    # without scratch an ensemble still gets its answer (single-CTA kernel), and the same one
    if w > 1:
        out = torch.empty_like(pair)
        hs = _native.handle_array([t.handle for t in trials])
        _native.check(_native.load().rfk_ensemble_forward(hs, w, w, x.data_ptr(), n, out.data_ptr(), None, 0, None))
        torch.cuda.synchronize()
        assert torch.equal(out, pair)
    for t in trials:
        t.close()


@pytest.mark.parametrize('n', [65536, 20000])
def test_fused_kernel_gives_the_same_bits_on_every_launch(n):
    """Row-independent predict (TfFeedForward.py:95) and a deterministic mean (ensemble.py:13-14):
    the same batch launched again and again, alone and as W identical trials (whose float64 mean is
    exact), must reproduce the first launch bit for bit.  (Round 1 failed this: two warps issued
    the MMAs of one accumulator.)"""
    dead = synth.dead_mask(784)
    p = synth.make_trial(3, 28, 1, 2, 128, dead=dead)
    x = torch.from_numpy(synth.make_queries(n, 28, 1, 'uniform', seed=5, dead=dead)).cuda()
    tr = engine.Trial(p)
    ref = tr.forward(x).clone()
    side = torch.cuda.Stream()
    junk = torch.empty(32 << 20, dtype=torch.uint8, device='cuda')
    launches = 0
    for w in (1, 2, 4, 8):
        ens = engine.Ensemble([tr] * w)
        for rep in range(6):
            if rep & 1:                       # vary the load: a copy stream runs beside every other launch
                with torch.cuda.stream(side):
                    junk.copy_(junk.flip(0))
            got = ens.forward(x)
            launches += 1
            assert torch.equal(got, ref), 'launch {} (w = {}) differs from the first'.format(launches, w)
    assert launches >= 20
    torch.cuda.synchronize()
    tr.close()


def test_full_size_properties_batch_65536():
    """BASELINE config 3 size: size-independent properties instead of an oracle pass."""
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, 2, 128, dead=dead) for t in range(8)]
    X = synth.make_queries(65536, 28, 1, 'uniform', seed=0, dead=dead)
    x = torch.from_numpy(X).cuda()
    trials = [engine.Trial(p) for p in params]
    out = engine.Ensemble(trials).forward(x)
    o = out.cpu().numpy()
    assert np.isfinite(o).all() and np.abs(o.sum(1) - 1).max() <= 1e-5
    # identical trials ensemble to the trial itself
    same = engine.Ensemble([trials[0]] * 4).forward(x)
    assert torch.equal(same, trials[0].forward(x))
    # rows are independent: any sub-batch gives the same rows, bit for bit
    part = engine.Ensemble(trials).forward(x[12800:12800 + 4096].contiguous())
    assert torch.equal(part, out[12800:12800 + 4096])
    perm = torch.randperm(65536, device='cuda')
    assert torch.equal(engine.Ensemble(trials).forward(x[perm].contiguous()), out[perm])
    # oracle on a bounded slice of the same batch
    sl = slice(57000, 57000 + 8192)
    want = ff.ensemble_mean_array([ff.forward(p, X[sl], np.float32) for p in params])
    assert np.abs(o[sl].astype(np.float64) - want).max() <= TOL_TC
    assert np.mean(o[sl].argmax(1) == want.argmax(1)) >= 0.999
    for t in trials:
        t.close()


def test_errors_are_reported_not_swallowed():
    params, X = _case(28, 1, 1, 8, 4)
    tr = engine.Trial(params)
    with pytest.raises(ValueError):
        tr.forward(torch.zeros((4, 100), dtype=torch.uint8, device='cuda'))
    with pytest.raises(TypeError):
        tr.forward(torch.zeros((4, 784), dtype=torch.float32, device='cuda'))
    other = engine.Trial(synth.make_trial(0, 32, 1, 1, 8))
    with pytest.raises(_native.RafikiNativeError):
        engine.Ensemble([tr, other]).forward(torch.from_numpy(X).cuda())
    bad = dict(params)
    bad['kernels'] = [params['kernels'][0]] + [np.zeros((9, 10), np.float32)]
    with pytest.raises(ValueError):
        engine.Trial(bad)


# ------------------------------------------------------------------------------- wide trials (layered tcgen05 GEMMs)
@pytest.mark.parametrize('S,ch,L,H,n', [(32, 3, 3, 2048, 300), (28, 1, 2, 300, 517), (32, 3, 1, 256, 129),
                                        (28, 1, 3, 64, 64)])
@pytest.mark.parametrize('compute,tol', [('f16', TOL_TC), ('bf16', TOL_BF16), ('f32', TOL_F32)])
def test_wide_trial_layered_path(S, ch, L, H, n, compute, tol):
    """BASELINE config 4 shape (D 3072, 3 x 2048) and other trials the fused kernel does not take."""
    params, X = _case(S, ch, L, H, n)
    tr = engine.Trial(params, compute=compute)
    assert not tr.fused
    got = _np(tr.forward(torch.from_numpy(X).cuda()))
    want = ff.forward(params, X, np.float32)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= tol
    assert np.abs(got.sum(1) - 1).max() <= 1e-5
    tr.close()


def test_wide_trial_in_an_ensemble_with_a_fused_one():
    dead = synth.dead_mask(784)
    params = [synth.make_trial(0, 28, 1, 2, 300, dead=dead), synth.make_trial(1, 28, 1, 2, 64, dead=dead)]
    X = synth.make_queries(700, 28, 1, 'uniform', seed=4, dead=dead)
    trials = [engine.Trial(p) for p in params]
    got = engine.Ensemble(trials).forward(torch.from_numpy(X).cuda()).cpu().numpy()
    want = ff.ensemble_mean_array([ff.forward(p, X, np.float32) for p in params])
    assert np.abs(got.astype(np.float64) - want).max() <= TOL_TC


def test_mixed_ensemble_with_an_external_plugin():
    """ensembles can mix model classes (examples/scripts/quickstart.py:78-82 trains
    TfFeedForward + SkDt); kernel 3 takes any [n, C] probability block, e.g. one uploaded from a
    CPU plugin's predict_proba."""
    params, X = _case(28, 1, 2, 64, 777)
    tr = engine.Trial(params)
    gpu_probs = tr.forward(torch.from_numpy(X).cuda())
    rng = np.random.default_rng(1)
    tree = rng.dirichlet(np.ones(10), size=777).astype(np.float32)          # a decision tree's predict_proba
    got = engine.ensemble_mean([gpu_probs, torch.from_numpy(tree).cuda()]).cpu().numpy()
    want = ff.ensemble_predictions([gpu_probs.cpu().numpy().tolist(), tree.tolist()], 'IMAGE_CLASSIFICATION')
    assert np.array_equal(got, np.asarray(want).astype(np.float32))
    tr.close()


def test_peer_reduce_single_rank_degenerates_to_a_rounded_copy():
    """The cross-GPU kernel with a world of one (the multi-rank case needs several GPUs and is
    checked by tests/test_gpu_multi.py under torchrun): flags, step counters and the slice
    arithmetic must work, and two steps in a row must reuse the flag block without a reset."""
    import ctypes
    lib = _native.load()
    n, c = 1000, 10
    flags = torch.zeros(3, dtype=torch.int32, device='cuda')
    for step in (1, 2):
        part = torch.rand((n, c), dtype=torch.float32, device='cuda')
        out = torch.zeros((n, c), dtype=torch.float32, device='cuda')
        probs = (ctypes.c_void_p * 1)(part.data_ptr())
        fl = (ctypes.c_void_p * 1)(flags.data_ptr())
        _native.check(lib.rfk_peer_enable(0))
        _native.check(lib.rfk_peer_reduce(probs, fl, 1, 0, 0, step, n * c, out.data_ptr(), 8, None))
        _native.check(lib.rfk_peer_wait(flags.data_ptr(), 1, step, None))
        torch.cuda.synchronize()
        assert torch.equal(out, part)
    assert flags.cpu().tolist() == [2, 2, 16]
