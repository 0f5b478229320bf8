"""Shard maps: the product's host logic against the oracle's restatement of the reference
(integer work: bit-exact), plus properties."""
import numpy as np


def test_table_owner_matches_reference_strategy(ffemb, oracle):
    for G in (1, 2, 4, 8):
        for t in range(26):
            assert ffemb.sharding.table_owner(t, G) == oracle.table_owner(t, G) == t % G
        owned = [ffemb.sharding.tables_of(g, 26, G) for g in range(G)]
        assert sorted(sum(owned, [])) == list(range(26))


def test_batch_block_matches_legion_blocked_partition(ffemb, oracle):
    for B in (1, 7, 256, 2048, 16384, 65537):
        for G in (1, 2, 3, 4, 8):
            blocks = [ffemb.sharding.batch_block(B, G, g) for g in range(G)]
            assert blocks == [oracle.batch_block(B, G, g) for g in range(G)]
            assert blocks[0][0] == 0 and blocks[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))  # disjoint + complete


def test_row_wise_maps(ffemb, oracle):
    rng = np.random.default_rng(0)
    for R, G in ((100, 8), (1000003, 8), (17, 4), (8, 8), (5, 8)):
        rows = rng.integers(0, R, 500)
        for r in rows:
            assert ffemb.sharding.row_owner(int(r), R, G) == oracle.lib().oracle_row_owner(int(r), R, G)
            assert ffemb.sharding.row_local(int(r), R, G) == oracle.lib().oracle_row_local(int(r), R, G)
        assert sum(ffemb.sharding.shard_rows(R, G, g) for g in range(G)) == R
        offsets, perm, local = oracle.route_rows(rows, R, G)
        # stable bucketing: positions ascending inside each bucket, owners correct
        for o in range(G):
            seg = perm[offsets[o]:offsets[o + 1]]
            assert np.all(np.diff(seg) > 0)
            assert np.all(rows[seg] // ffemb.sharding.rows_per_shard(R, G) == o)
            assert np.array_equal(local[offsets[o]:offsets[o + 1]], rows[seg] % ffemb.sharding.rows_per_shard(R, G))


def test_exchange_counts_are_transposes(ffemb):
    for G in (1, 2, 4, 8):
        plan = ffemb.sharding.TableWisePlan(26, G, 16384 + 3, 16)
        send = [plan.send_counts(g) for g in range(G)]
        recv = [plan.recv_counts(g) for g in range(G)]
        for g in range(G):
            for h in range(G):
                assert send[g][h] == recv[h][g]
        lo_hi = [ffemb.sharding.batch_block(plan.batch, G, g) for g in range(G)]
        assert sum(sum(r) for r in recv) == 26 * plan.batch * 16
        assert all(sum(recv[g]) == 26 * (lo_hi[g][1] - lo_hi[g][0]) * 16 for g in range(G))


def test_concat_layout(oracle):
    """Row b of the interaction input = [x(b), e_0(b), ..., e_{T-1}(b)] (concat.cu:211-251)."""
    rng = np.random.default_rng(1)
    ins = [rng.uniform(-1, 1, (5, w)).astype(np.float32) for w in (3, 4, 4, 2)]
    out = oracle.concat_forward(ins)
    assert np.array_equal(out, np.concatenate(ins, axis=1))
    grads = oracle.concat_backward(out, [3, 4, 4, 2])
    for a, b in zip(grads, ins):
        assert np.array_equal(a, b)


def test_synth_configs(ffemb):
    for n in range(1, 6):
        wl = ffemb.synth.config(n)
        assert wl.num_tables == len(wl.bags)
    wl = ffemb.synth.config(2)
    assert wl.num_tables == 26 and wl.dim == 16 and wl.batch == 16384 and wl.lookups == 425984
    assert max(wl.rows) == 10131227
    assert ffemb.synth.config(3).lookups > 10_000_000
    small = ffemb.synth.config(5, scale=0.001)
    idx = ffemb.synth.gen_indices(small, 0)
    assert idx.shape == (small.batch, 64) and idx.min() >= 0 and idx.max() < small.rows[0]
    # Zipf: the hottest row takes a large share
    top = np.bincount(idx.reshape(-1)).max() / idx.size
    assert top > 0.02
    assert np.array_equal(idx, ffemb.synth.gen_indices(small, 0))  # seeded


# ---- strategy files: the reference's text format -------------------------------
def _golden_strategy():
    import os
    return os.path.join(os.path.dirname(__file__), "golden", "dlrm_tablewise_26x8.strategy")


def test_strategy_reader_parses_the_reference_writers_file(ffemb):
    """tests/golden/dlrm_tablewise_26x8.strategy was written by the reference's own
    save_strategies_to_file (tests/golden/make_strategy_golden.py): 26 tables, table i whole on
    GPU i % 8, concat data-parallel."""
    S = ffemb.sharding
    st = S.read_strategy(_golden_strategy())
    assert len(st) == 27
    assert S.owners_from_strategy(st, 26) == [i % 8 for i in range(26)] == S.place_tables([1] * 26, [1] * 26, 128, 8, "modulo")
    assert st["concat"].dims == [1, 8] and st["concat"].device_ids == list(range(8))
    assert all(st[f"embedding{i}"].dims == [1, 1] and st[f"embedding{i}"].device_type == S.DEVICE_GPU for i in range(26))


def test_strategy_writer_emits_the_reference_writers_bytes(ffemb, tmp_path):
    S = ffemb.sharding
    p = str(tmp_path / "s.strategy")
    S.write_strategy(p, S.tablewise_strategy([i % 8 for i in range(26)], 8))
    assert open(p).read() == open(_golden_strategy()).read()
    assert S.read_strategy(p) == S.read_strategy(_golden_strategy())


def test_strategy_writer_is_read_by_the_reference_loader(ffemb, oracle, tmp_path):
    """A cost-balanced placement written by the product, parsed by the reference's own
    load_strategies_from_file (oracle/_ref, src/runtime/strategy.cc:95-146)."""
    import pytest
    if not oracle.ref_cpu_available():
        pytest.skip("oracle/_ref/libffref_cpu.so not built (needs /root/reference)")
    S = ffemb.sharding
    wl = ffemb.synth.config(3)
    own = S.place_tables(wl.rows, wl.bags, wl.dim, 8, "balanced")
    assert sorted(set(own)) == list(range(8))
    p = str(tmp_path / "balanced.strategy")
    S.write_strategy(p, S.tablewise_strategy(own, 8))
    for i, g in enumerate(own):
        assert oracle.ref_strategy_load_op(p, f"embedding{i}") == (S.DEVICE_GPU, [1, 1], [g])
    assert oracle.ref_strategy_load_op(p, "concat") == (S.DEVICE_GPU, [1, 8], list(range(8)))
    assert S.owners_from_strategy(S.read_strategy(p), wl.num_tables) == own


def test_strategy_errors(ffemb, tmp_path):
    import pytest
    S = ffemb.sharding
    p = tmp_path / "bad.strategy"
    p.write_text("1\nembedding0\n7\n2\n1\t1\t\n1\n0\t\n")          # unknown device type
    with pytest.raises(ValueError):
        S.read_strategy(str(p))
    p.write_text("1\nembedding0\n0\n2\n1\t2\t\n1\n0\t\n")          # 2 parts, 1 device id
    with pytest.raises(ValueError):
        S.read_strategy(str(p))
    p.write_text("1\nembedding0\n0\n2\n1\t2\t\n2\n0\t1\t\n")       # a table split over two GPUs
    with pytest.raises(ValueError):
        S.owners_from_strategy(S.read_strategy(str(p)), 1)
    with pytest.raises(KeyError):
        S.owners_from_strategy({}, 1)


# ---- shard plans with column splits and pipeline groups (sharding.plan_shards) -------------------
def _check_plan(S, parts, rows, dim, world):
    # every column of every table exactly once; parts of a split table on distinct ranks
    for t in range(len(rows)):
        ps = sorted([p for p in parts if p.table == t], key=lambda p: p.col0)
        assert ps and ps[0].col0 == 0 and sum(p.width for p in ps) == dim
        for a, b in zip(ps, ps[1:]):
            assert a.col0 + a.width == b.col0
        assert len({p.rank for p in ps}) == len(ps)
        assert all(p.width % 4 == 0 and 0 <= p.rank < world for p in ps)
    # the kernels take one row width per launch: a (rank, group) never mixes widths
    for g in range(world):
        for k in range(S.num_groups(parts)):
            assert len({p.width for p in parts if p.rank == g and p.group == k}) <= 1


def test_plan_modulo_is_the_reference_placement(ffemb):
    S, wl = ffemb.sharding, ffemb.synth.config(3)
    for world in (1, 2, 8):
        parts = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "modulo", split=True, groups=1)
        assert [(p.table, p.col0, p.width, p.rank, p.group) for p in parts] == \
            [(t, 0, wl.dim, t % world, 0) for t in range(wl.num_tables)]


def test_plan_balanced_splits_the_heaviest_table_and_levels_the_ranks(ffemb):
    S, wl = ffemb.sharding, ffemb.synth.config(3)
    cost = lambda ps: sum(S.part_cost(wl.rows[p.table], wl.bags[p.table], wl.dim, p.width) for p in ps)
    for world in (2, 4, 8):
        for groups in (1, 2):
            parts = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "balanced", split=True, groups=groups)
            again = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "balanced", split=True, groups=groups)
            assert parts == again  # deterministic: every rank derives the same plan
            _check_plan(S, parts, wl.rows, wl.dim, world)
            loads = [cost([p for p in parts if p.rank == g]) for g in range(world)]
            assert max(loads) <= 1.15 * (sum(loads) / world)
            modulo = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "modulo")
            assert max(loads) <= max(cost([p for p in modulo if p.rank == g]) for g in range(world))
    parts8 = S.plan_shards(wl.rows, wl.bags, wl.dim, 8, "balanced", split=True, groups=2)
    assert sorted((p.col0, p.width) for p in parts8 if p.table == 0) == [(0, 64), (64, 64)]  # 40M rows x bag 20
    whole = S.plan_shards(wl.rows, wl.bags, wl.dim, 8, "balanced", split=False, groups=2)
    assert all(p.width == wl.dim for p in whole)
    for world in (2, 4):  # per-rank share above the heaviest table: nothing to split
        assert all(p.width == wl.dim for p in S.plan_shards(wl.rows, wl.bags, wl.dim, world, "balanced", True, 2))


def test_plan_round_trips_through_the_strategy_file_format(ffemb, tmp_path):
    S, wl = ffemb.sharding, ffemb.synth.config(3)
    parts = S.plan_shards(wl.rows, wl.bags, wl.dim, 8, "balanced", split=True, groups=1)
    path = str(tmp_path / "plan.strategy")
    S.write_strategy(path, S.plan_strategy(parts, 8))
    back = S.parts_from_strategy(S.read_strategy(path), wl.num_tables, wl.dim)
    key = lambda p: (p.table, p.col0, p.width, p.rank)
    assert sorted(map(key, back)) == sorted(map(key, parts))
    # a plan without splits is a plain table-wise strategy of the reference's kind: [1,1] grids
    whole = S.plan_shards(wl.rows, wl.bags, wl.dim, 8, "balanced", split=False)
    st = S.plan_strategy(whole, 8)
    assert S.owners_from_strategy(st, wl.num_tables) == [next(p.rank for p in whole if p.table == t)
                                                         for t in range(wl.num_tables)]


# ---- row-wise routed backward: the host statement of the slot arrays (include/ffemb.h, ffemb_rowwise_route)
def test_route_slots_is_the_oracles_stable_bucketing(ffemb, oracle):
    S = ffemb.sharding
    rng = np.random.default_rng(11)
    for R, G, B, L in [(1000, 4, 37, 3), (17, 8, 64, 1), (5, 2, 9, 7), (4096, 3, 50, 4)]:
        idx = rng.integers(0, R, (B, L), dtype=np.int64)
        offsets, perm, local = oracle.route_rows(idx.reshape(-1), R, G)
        for g in range(G):
            slots = S.route_slots(idx, R, G, g)
            assert slots.shape == idx.shape and slots.dtype == np.int64
            pos = np.flatnonzero(slots.reshape(-1) >= 0)
            # same positions, same order (global (sample, j) order) and same local rows as the
            # oracle's stable bucketing; everything else is an empty slot
            lo, hi = offsets[g], offsets[g + 1]
            assert np.array_equal(pos, perm[lo:hi])
            assert np.array_equal(slots.reshape(-1)[pos], local[lo:hi])
            assert np.all(slots.reshape(-1)[np.setdiff1d(np.arange(B * L), pos)] == -1)
            assert slots.max(initial=-1) < S.shard_rows(R, G, g)
        # every lookup lands on exactly one owner
        owners = sum((S.route_slots(idx, R, G, g) >= 0).astype(np.int64) for g in range(G))
        assert np.all(owners == 1)


def test_routed_backward_on_slot_arrays_equals_the_unsharded_step(ffemb, oracle):
    """Owner g's backward over its slot array (empty slots skipped, equal rows met in global
    (sample, j) order) + SGD on its shard == the rows of the unsharded reference step, bit for bit."""
    S = ffemb.sharding
    rng = np.random.default_rng(12)
    for R, G, B, L, D, aggr in [(300, 4, 48, 3, 8, oracle.AGGR_SUM), (64, 8, 128, 2, 16, oracle.AGGR_SUM),
                                (33, 2, 40, 5, 4, oracle.AGGR_AVG)]:
        idx = rng.integers(0, R, (B, L), dtype=np.int64)
        idx[: B // 4] = idx[0]  # heavy duplicates: order of accumulation matters
        W = rng.standard_normal((R, D)).astype(np.float32)
        og = rng.standard_normal((B, D)).astype(np.float32)
        want = oracle.reference_step(idx, og, W, aggr=aggr, lr=0.05)
        rps = S.rows_per_shard(R, G)
        got = np.empty_like(W)
        for g in range(G):
            n = S.shard_rows(R, G, g)
            if n == 0:
                continue
            slots = S.route_slots(idx, R, G, g)
            # an empty slot touches nothing: send it to a throw-away row past the shard
            filled = np.where(slots >= 0, slots, n)
            grad = oracle.backward_dense(filled, og, n + 1, aggr=aggr)[:n]
            got[g * rps:g * rps + n] = oracle.sgd_update(W[g * rps:g * rps + n], grad, 0.05)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_plan_groups_are_consistent_across_ranks(ffemb):
    """Every rank derives the plan alone (no exchange of plans): same inputs -> same parts, and the
    pipeline groups partition every rank's parts — num_groups(plan) is a property of the PLAN, not of
    one rank's share (the world-8 IndexError of round 2)."""
    S, wl = ffemb.sharding, ffemb.synth.config(3)
    for world in (2, 4, 8):
        for groups in (1, 2, 3):
            a = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "balanced", True, groups)
            b = S.plan_shards(wl.rows, wl.bags, wl.dim, world, "balanced", True, groups)
            key = lambda p: (p.table, p.col0, p.width, p.rank, p.group)
            assert list(map(key, a)) == list(map(key, b))
            K = S.num_groups(a)
            assert K >= 1
            assert all(0 <= p.group < K for p in a)
            for r in range(world):
                mine = [p for p in a if p.rank == r]
                assert max((p.group for p in mine), default=0) < K


def test_route_bytes_accounting(ffemb):
    S = ffemb.sharding
    # one table, bag 4, world 2, batch 8: ext 4 samples; slots 4*4*8*2 = 256 B; rows 4 bags * 2 owners * 16*4
    assert S.route_bytes(8, [4], 16, 2) == 256 + 4 * 2 * 64
    # bag 1: one owner per bag whatever the world
    assert S.route_bytes(8, [1], 16, 8) == 1 * 1 * 8 * 8 + 1 * 1 * 64
    assert S.route_bytes(8, [4], 16, 2, owners_per_bag=1.5) == 256 + int(4 * 1.5 * 64)
    # grows with the world on the slot side only
    assert S.route_bytes(1024, [2, 2], 64, 8) > S.route_bytes(1024, [2, 2], 64, 2) // 4


# ---- property tests (hypothesis): strategy files and shard plans over random shapes ---------------
def test_random_strategy_files_round_trip_and_load_in_the_reference(ffemb, oracle, tmp_path):
    import pytest
    from hypothesis import given, settings, strategies as st, HealthCheck
    S = ffemb.sharding
    have_ref = oracle.ref_cpu_available()

    name = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789_", min_size=1, max_size=24)
    dims = st.lists(st.integers(1, 4), min_size=1, max_size=4).filter(lambda d: int(np.prod(d)) <= 32)

    @st.composite
    def config(draw):
        d = draw(dims)
        n = int(np.prod(d))
        ids = draw(st.lists(st.integers(0, 1023), min_size=n, max_size=n))
        return S.ParallelConfig(draw(st.sampled_from([S.DEVICE_GPU, S.DEVICE_CPU])), d, ids)

    case = {"n": 0}

    @settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.function_scoped_fixture])
    @given(st.dictionaries(name, config(), min_size=1, max_size=12))
    def run(strategies):
        case["n"] += 1
        p = str(tmp_path / f"s{case['n']}.strategy")
        S.write_strategy(p, strategies)
        back = S.read_strategy(p)
        assert back == strategies
        if have_ref:  # the reference's own loader sees the same thing, op by op
            for op, c in strategies.items():
                assert oracle.ref_strategy_load_op(p, op) == (c.device_type, c.dims, c.device_ids)

    run()
    assert case["n"] >= 10


def test_random_shard_plans_cover_every_column_once_and_survive_the_file_format(ffemb, tmp_path):
    from hypothesis import given, settings, strategies as st, HealthCheck
    S = ffemb.sharding
    case = {"n": 0}

    @settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.function_scoped_fixture])
    @given(st.lists(st.tuples(st.integers(1, 50_000_000), st.integers(1, 64)), min_size=1, max_size=30),
           st.sampled_from([16, 32, 64, 128, 256]), st.sampled_from([1, 2, 3, 4, 8]),
           st.sampled_from(["balanced", "modulo"]), st.booleans(), st.integers(1, 3))
    def run(tables, dim, world, kind, split, groups):
        case["n"] += 1
        rows, bags = [r for r, _ in tables], [l for _, l in tables]
        parts = S.plan_shards(rows, bags, dim, world, kind, split, groups)
        _check_plan(S, parts, rows, dim, world)
        assert all(0 <= p.group < S.num_groups(parts) for p in parts)
        if not split or kind == "modulo":
            assert all(p.width == dim for p in parts)
        if kind == "modulo":
            assert [p.rank for p in sorted(parts, key=lambda p: p.table)] == [t % world for t in range(len(rows))]
        path = str(tmp_path / f"p{case['n']}.strategy")
        S.write_strategy(path, S.plan_strategy(parts, world))
        back = S.parts_from_strategy(S.read_strategy(path), len(rows), dim)
        key = lambda p: (p.table, p.col0, p.width, p.rank)
        assert sorted(map(key, back)) == sorted(map(key, parts))

    run()
    assert case["n"] >= 10


def test_plan_never_cuts_a_table_into_more_parts_than_ranks_and_files_hold_uneven_cuts(ffemb, tmp_path):
    """Found by the property test above: one heavy table on 2 ranks used to be cut into quarters, two of
    them on each rank (the unsplit half again, with a second index sort); on 3 ranks the cut is a half
    next to two quarters, which the [c,1] grid of the file holds as 4 cells a a b c."""
    S = ffemb.sharding
    two = S.plan_shards([1], [1], 64, 2, "balanced", True, 1)
    assert sorted((p.col0, p.width, p.rank) for p in two) == [(0, 32, 0), (32, 32, 1)]
    three = S.plan_shards([1], [1], 64, 3, "balanced", True, 1)
    assert sorted(p.width for p in three) == [16, 16, 32] and len({p.rank for p in three}) == 3
    st = S.plan_strategy(three, 3)
    assert st["embedding0"].dims == [4, 1] and len(set(st["embedding0"].device_ids)) == 3
    path = str(tmp_path / "uneven.strategy")
    S.write_strategy(path, st)
    key = lambda p: (p.table, p.col0, p.width, p.rank)
    assert sorted(map(key, S.parts_from_strategy(S.read_strategy(path), 1, 64))) == sorted(map(key, three))
