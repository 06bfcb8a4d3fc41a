"""Failure detection and state handling of the fused round kernel (csrc/round.cu), on emulated ranks (all ranks of
the protocol in ONE cooperative launch on one GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _lib(native_lib):
    return native_lib


def _mk(G, counts, P, seed):
    torch.manual_seed(seed)
    g = torch.randn(P, device=DEV)
    clients = [(g + 0.1 * torch.randn(c, P, device=DEV)) if c else None for c in counts]
    weights = [(torch.randint(10, 100, (c,), device=DEV).float()) if c else None for c in counts]
    total = sum(float(w.sum()) for w in weights if w is not None)
    return g, clients, weights, total


def test_clipped_round_after_a_round_where_one_rank_had_no_clients():
    """A rank that drew no client skips the intra-rank grid barrier of the clip phase; its barrier target must not run
    ahead of its counter (the next clipped round with clients would wait forever)."""
    from fedml_b200.parallel.round_engine import DPConfig, RoundEngine, emulated_round
    from test_gpu_kernels import _oracle_round
    RoundEngine.set_spin_timeout(5000)
    P, P_train = 4 * 3001, 4 * 2800
    dpc = DPConfig(clip=1.5, sigma=0.0, seed=5)
    state = None
    g_cur = None
    for rnd, counts in enumerate([(3, 0), (2, 4), (0, 3), (5, 1)]):
        g, clients, weights, total = _mk(2, counts, P, 20 + rnd)
        if g_cur is None:
            g_cur = g
        else:                                       # continue from the model the previous round left on the ranks
            clients = [None if c is None else c - g + g_cur for c in clients]
        outs, state = emulated_round(clients, weights, g_cur, P_train, 1.0 / total, False, dp=dpc, round_idx=rnd,
                                     epoch=rnd + 1, state=state)
        RoundEngine.check_spin_errors()
        ref = _oracle_round(clients, weights, g_cur, P_train, dpc.clip, 0.0, dpc.seed, rnd, None, None, None, 1, 1.0 / total)
        torch.testing.assert_close(outs[0].double(), ref, rtol=1e-5, atol=2e-5)
        assert torch.equal(outs[0], outs[1])
        g_cur = outs[0].clone()
    RoundEngine.set_spin_timeout(20000)


@pytest.mark.parametrize("opt_name,momentum", [("adam", 0.0), ("sgd", 0.9)])
def test_two_shot_sharded_optimizer_state_resume(opt_name, momentum):
    """Two-shot mode keeps valid moments for shard r on rank r only.  A checkpoint assembled from every rank's own shard,
    loaded on all ranks, continues exactly like the uninterrupted run."""
    from fedml_b200.parallel.round_engine import ServerOptConfig, emulated_round, merge_sharded_moments
    G, P, P_train, K = 2, 4 * 2048, 4 * 2000, 8
    opt = ServerOptConfig(opt_name, lr=0.05, momentum=momentum, beta1=0.9, beta2=0.99, eps=1e-3)
    torch.manual_seed(3)
    g0 = torch.randn(P, device=DEV)

    def round_inputs(rnd, g):
        torch.manual_seed(100 + rnd)
        X = g + 0.05 * torch.randn(K, P, device=DEV)
        w = torch.arange(1, K + 1, device=DEV).float()
        per = K // G
        return [X[r * per:(r + 1) * per] for r in range(G)], [w[r * per:(r + 1) * per].contiguous() for r in range(G)], float(w.sum())

    def run(state, g, rnd):
        cl, wt, tot = round_inputs(rnd, g)
        outs, state = emulated_round(cl, wt, g, P_train, 1.0 / tot, True, opt=opt, opt_step=rnd + 1, epoch=rnd + 1, state=state)
        return outs[0].clone(), state

    # uninterrupted: three rounds
    g, st = g0.clone(), None
    for rnd in range(3):
        g, st = run(st, g, rnd)
    full = g
    # interrupted after two rounds: checkpoint = global + moments merged from each rank's own shard
    g, st = g0.clone(), None
    for rnd in range(2):
        g, st = run(st, g, rnd)
    m = merge_sharded_moments(st["m"], P, P_train) if st["m"][0] is not None else None
    v = merge_sharded_moments(st["v"], P, P_train) if st["v"][0] is not None else None
    # the shards a rank does not own really are stale (otherwise this test checks nothing)
    if m is not None:
        assert not torch.equal(st["m"][0], st["m"][1])
    cl, wt, tot = round_inputs(2, g)
    outs, _ = emulated_round(cl, wt, g, P_train, 1.0 / tot, True, opt=opt, m=m, v=v, opt_step=3, epoch=1, state=None)
    assert torch.equal(outs[0], full) and torch.equal(outs[1], full)
    # resuming from rank 0's copy alone (the old behaviour) does NOT reproduce the run
    m0 = st["m"][0] if m is not None else None
    v0 = st["v"][0] if v is not None else None
    outs_bad, _ = emulated_round(cl, wt, g, P_train, 1.0 / tot, True, opt=opt, m=m0, v=v0, opt_step=3, epoch=1, state=None)
    assert not torch.equal(outs_bad[1], full)


def test_dead_rank_is_reported_not_waited_for_forever():
    """Fault injection: one emulated rank never announces the round.  The other ranks' waits give up after the
    time-out, the kernel retires, the host gets FB_ERR_TIMEOUT naming the rank, and the next round works."""
    import time
    from fedml_b200.ops.native import NativeError
    from fedml_b200.parallel.round_engine import RoundEngine, emulated_round
    P, P_train = 4 * 3001, 4 * 2800
    g, clients, weights, total = _mk(3, (2, 3, 1), P, 1)
    RoundEngine.set_spin_timeout(300)
    try:
        t0 = time.time()
        outs, state = emulated_round(clients, weights, g, P_train, 1.0 / total, False, epoch=1, dead_rank=1)
        torch.cuda.synchronize()
        assert time.time() - t0 < 30.0
        with pytest.raises(NativeError, match="timed out"):
            RoundEngine.check_spin_errors()
        RoundEngine.check_spin_errors()                      # record was cleared
        # healthy round afterwards, same flag buffers (epoch moves on)
        for r in range(3):
            state["globals"][r].copy_(g)
        outs, state = emulated_round(clients, weights, g, P_train, 1.0 / total, False, epoch=2, state=state)
        RoundEngine.check_spin_errors()
        ref = sum((w[:, None] * c).sum(0) for w, c in zip(weights, clients)) / total
        torch.testing.assert_close(outs[0], ref, rtol=1e-5, atol=2e-5)
    finally:
        RoundEngine.set_spin_timeout(20000)


@pytest.mark.parametrize("K,P", [(512, 4 * 5003), (300, 4 * 1001), (192, 2048)])
def test_many_client_median_and_trimmed_mean_sorted_runs(K, P):
    """K > 128 rows: in-place sorted runs + merge-select (fb_robust_sort_runs / fb_robust_select_runs) == torch (lower
    median, trimmed mean), incl. +inf padding rows that must not count."""
    from fedml_b200 import ops
    torch.manual_seed(K)
    x = torch.randn(K, P, device=DEV)
    x[:, :64] = torch.randint(-2, 3, (K, 64), device=DEV).float()        # heavy ties
    x[:, 64:96:3] = 5.0            # coordinates where every client holds the same value, interleaved with ones that differ
    med = ops.coordinate_median(x)
    assert torch.equal(med, torch.median(x, dim=0).values)
    trim = K // 10
    tm = ops.trimmed_mean(x, trim)
    ref = torch.sort(x, dim=0).values[trim:K - trim].double().mean(dim=0)
    torch.testing.assert_close(tm.double(), ref, rtol=1e-5, atol=1e-6)
    # padded layout as the multi-GPU engine uses it: G blocks of C_max rows, the tail of each block is +inf
    G, C_max = 4, (K + 3) // 4 + 3
    L = ops.run_length_for(C_max)
    if L:
        pad = torch.full((G * C_max, P), float("inf"), device=DEV)
        counts = [(K + G - 1 - g) // G for g in range(G)]
        o = 0
        for g, n in enumerate(counts):
            pad[g * C_max:g * C_max + n] = x[o:o + n]
            o += n
        rows = ops.ptr_table([pad.data_ptr() + k * P * 4 for k in range(G * C_max)], DEV)
        ops.robust_sort_runs(rows, G * C_max, L, 0, P)
        out = torch.empty(P, device=DEV)
        outs = ops.ptr_table([out.data_ptr()], DEV)
        ops.robust_select_runs(rows, G * C_max, K, L, 0, P, "median", 0, outs, 1)
        assert torch.equal(out, torch.median(x, dim=0).values)
        ops.robust_select_runs(rows, G * C_max, K, L, 0, P, "trimmed_mean", trim, outs, 1)
        torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-6)


def test_nccl_simulator_secure_aggregation_on_the_field_kernels():
    """secure_aggregation in the NCCL simulator on the GPU: fb_ff_quantize / fb_ff_mask (round nonce) / fb_ff_add /
    fb_ff_dequantize give the plain FedAvg model up to the quantisation step."""
    import fedml_b200 as fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.ops import native
    from fedml_b200.simulation import SimulatorNCCL
    outs = []
    for sec in (False, True):
        a = fedml.init(Arguments.from_dict(dict(
            training_type="simulation", backend="NCCL", dataset="mnist", model="lr", client_num_per_round=6, comm_round=2,
            batch_size=10, learning_rate=0.03, frequency_of_the_test=0, synthetic_train_size=1200, mnist_client_num=12,
            log_level="ERROR", using_gpu=True, compute_dtype="fp32", secure_aggregation=sec, random_seed=0)))
        ds, ncls = fedml.data.load(a)
        torch.manual_seed(0)
        sim = SimulatorNCCL(a, torch.device(DEV), ds, fedml.models.create(a, ncls)).simulator
        before = native.LAUNCHES[0]
        for r in range(2):
            sim.train_one_round(r)
        torch.cuda.synchronize()
        outs.append((sim.global_flat.clone(), native.LAUNCHES[0] - before))
    assert float((outs[0][0] - outs[1][0]).abs().max()) < 6 * 4 / 2 ** 16
    assert outs[1][1] > outs[0][1]                       # the field kernels ran


@pytest.mark.parametrize("two_shot", [False, True])
def test_flag_protocol_is_insensitive_to_block_and_rank_skew(two_shot):
    """Race-detection stand-in (compute-sanitizer is closed on this pool): every (rank, block, handshake) of the emulated
    4-rank round stalls a pseudo-random time below 150 us before it signals, with five different seeds, for three rounds in a
    row on the same buffers; the result must be bit-identical to the run without delays and identical on every rank."""
    import ctypes as C
    from fedml_b200.ops.native import lib
    from fedml_b200.parallel.round_engine import DPConfig, RoundEngine, ServerOptConfig, emulated_round
    RoundEngine.set_spin_timeout(5000)
    G, P, P_train = 4, 4 * 50021, 4 * 50000
    opt = ServerOptConfig("adam", lr=0.01)
    dpc = DPConfig(clip=2.0, sigma=0.001, seed=3)

    def run(seed, max_ns):
        assert lib().fb_round_set_skew(C.c_uint32(seed), C.c_uint32(max_ns)) == 0
        state, g_cur, outs_all = None, None, []
        for rnd in range(3):
            g, clients, weights, total = _mk(G, (3, 1, 4, 2), P, 40 + rnd)
            if g_cur is None:
                g_cur = g
            else:
                clients = [c - g + g_cur for c in clients]
            outs, state = emulated_round(clients, weights, g_cur, P_train, 1.0 / total, two_shot, opt=opt, dp=dpc,
                                         opt_step=rnd + 1, round_idx=rnd, epoch=rnd + 1, state=state)
            RoundEngine.check_spin_errors()
            for r in range(1, G):
                assert torch.equal(outs[0], outs[r])
            g_cur = outs[0].clone()
            outs_all.append(g_cur)
        return outs_all

    try:
        base = run(0, 0)
        for seed in range(1, 6):
            skewed = run(seed, 150_000)
            for a, b in zip(base, skewed):
                assert torch.equal(a, b), f"seed {seed}: result depends on the interleaving"
    finally:
        lib().fb_round_set_skew(C.c_uint32(0), C.c_uint32(0))
        RoundEngine.set_spin_timeout(20000)
