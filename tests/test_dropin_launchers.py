"""CPU test of the LAUNCHER side of the fast path behind Legion:
ffemb_group_forward / ffemb_group_prepare / ffemb_group_backward_sgd of
dlrm-flexflow_b200/csrc/flexflow/embedding_group_ffemb.cu — what replaces Embedding::forward() / backward()
(/root/reference/src/ops/embedding.cu:274-306, 356-388) for the tables of one GPU — compiled against the
reference's own headers and run against a Legion stand-in that writes the index launch down
(oracle/shim/legion.h: Runtime::index_launch_hook; oracle/ref_glue_gpu.cu: ffref_launch_record) — and, the same
way, the reference's own Embedding::forward() / backward() from the reference's object code, side by side.

Nothing is launched and no device is touched (the pointers below are just numbers).  What is held: the task
id, the single launch point of the table-wise [1,1] grid, the by-value arguments (table count, learning rate
taken from the SGD optimizer, prepared flag), the per-point OpMeta*, and per table the region requirements
in the order, privilege and field the leaf bodies read them — and that the (pointer, rectangle) frame made
from those requirements is the very frame the GPU tests hand the leaf bodies through
ffdropin_gpu_group_task (test_dropin_group_leaf_tasks_match_reference_per_table_tasks)."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def rec(oracle):
    if not oracle.dropin_gpu_available():
        pytest.skip("oracle/_ref/libffdropin_gpu.so not built (needs /root/reference)")
    try:
        oracle.dropin_gpu()
    except OSError as e:  # CUDA runtime libraries absent on this host
        pytest.skip(f"libffdropin_gpu.so does not load here: {e}")
    return lambda *a, **k: oracle.launch_record("fused", *a, **k)


BAGS, ROWS, D, B = [3, 1, 20, 7], [1000, 7, 40_000_000, 155], 128, 65536
IDX = [0x10000 * (t + 1) for t in range(4)]
OUT = [0x20000000 + 0x10000 * t for t in range(4)]
TAB = [0x7F0000000000 + 0x100000000 * t for t in range(4)]


def _common(r, n):
    c = r["consts"]
    assert r["points"] == 1 and r["point_args"] == 1          # one point: the [1,1] grid of the table-wise strategy
    assert r["local_arg_is_lead_meta"] and r["local_arglen"] == 8
    assert not r["must"] and r["mapper_id"] == 0
    assert r["args"]["len"] == 16 and r["args"]["num_tables"] == n and r["args"]["parity"] == 0
    for q in r["reqs"]:
        assert q["coherence"] == c["EXCLUSIVE"] and q["fields"] == [c["FID_DATA"]]
        # the reference maps the embedding output / output-grad to zero-copy host memory
        # (src/ops/embedding.cu:296-298, 378-380); the fused mode keeps them in framebuffer memory
        assert q["tag"] != c["MAP_TO_ZC_MEMORY"] and q["tag"] == 0
    assert r["tag"] != 0  # the lead op's mapping tag (hash of its name)
    return c


def test_group_forward_launcher(rec):
    r = rec(0, IDX, OUT, TAB, BAGS, D, B, ROWS)
    c = _common(r, 4)
    assert r["task_id"] == c["TASK_FORWARD"] and len(r["reqs"]) == 12
    for t in range(4):
        i, o, w = r["reqs"][3 * t:3 * t + 3]
        assert (i["privilege"], i["base"], i["extent"]) == (c["READ_ONLY"], IDX[t], (BAGS[t], B))
        assert (o["privilege"], o["base"], o["extent"]) == (c["WRITE_ONLY"], OUT[t], (D, B))
        assert (w["privilege"], w["base"], w["extent"]) == (c["READ_ONLY"], TAB[t], (ROWS[t], D))


def test_group_prepare_launcher_lists_indices_and_tables_only(rec):
    r = rec(1, IDX, None, TAB, BAGS, D, B, ROWS)
    c = _common(r, 4)
    assert r["task_id"] == c["TASK_PREPARE"] and len(r["reqs"]) == 8
    for t in range(4):
        i, w = r["reqs"][2 * t:2 * t + 2]
        assert (i["privilege"], i["base"], i["extent"]) == (c["READ_ONLY"], IDX[t], (BAGS[t], B))
        assert (w["privilege"], w["base"], w["extent"]) == (c["READ_ONLY"], TAB[t], (ROWS[t], D))


@pytest.mark.parametrize("prepared", [False, True])
def test_group_backward_launcher_updates_the_table_in_place(rec, prepared):
    r = rec(2, IDX, OUT, TAB, BAGS, D, B, ROWS, lr=0.01, prepared=prepared)
    c = _common(r, 4)
    assert r["task_id"] == c["TASK_BACKWARD_SGD"] and len(r["reqs"]) == 12
    # SGDOptimizer::lr is a double (include/optimizer.h:55); the task gets it as the float sgd_update takes
    assert r["args"]["lr"] == float(np.float32(0.01)) and r["args"]["prepared"] == int(prepared)
    for t in range(4):
        i, g, w = r["reqs"][3 * t:3 * t + 3]
        assert (i["privilege"], i["base"], i["extent"]) == (c["READ_ONLY"], IDX[t], (BAGS[t], B))
        assert (g["privilege"], g["base"], g["extent"]) == (c["READ_ONLY"], OUT[t], (D, B))
        # region 2 is the TABLE, read-write — the reference lists the weight-GRAD here
        # (src/ops/embedding.cu:384-385) and leaves the update to SGDOptimizer::update
        assert (w["privilege"], w["base"], w["extent"]) == (c["READ_WRITE"], TAB[t], (ROWS[t], D))


def test_one_table_groups_and_task_ids_are_the_custom_gpu_ids(rec):
    r = rec(0, IDX[:1], OUT[:1], TAB[:1], BAGS[:1], 16, 7, ROWS[:1])
    c = _common(r, 1)
    assert len(r["reqs"]) == 3 and r["reqs"][1]["extent"] == (16, 7)
    # include/model.h keeps CUSTOM_GPU_TASK_ID_1.. free for this; three distinct consecutive ids
    assert c["TASK_PREPARE"] == c["TASK_FORWARD"] + 1 and c["TASK_BACKWARD_SGD"] == c["TASK_FORWARD"] + 2


def test_fused_create_weights_differs_from_the_reference_in_exactly_two_arguments(oracle, rec):
    """Embedding::create_weights of the reference (src/ops/embedding.cu:58-74, run here from the reference's
    own object code) against ffemb_fused_create_weights: same task space (asked for by the op's name, 2-D),
    same {out_channels, num_entries} fp32 weight with the op's initializer — but no gradient region and no
    parameter synchronisation, so neither Op::zero_grad nor SGDOptimizer::update nor the gradient
    all-reduce ever sees the table."""
    if not oracle.ref_gpu_available():
        pytest.skip("oracle/_ref/libffref_gpu.so not built")
    ref = oracle.create_weights_record("reference", "embedding7", rows=39_884_406, dim=128)
    ours = oracle.create_weights_record("fused", "embedding7", rows=39_884_406, dim=128)
    c = ref["consts"]
    for r in (ref, ours):
        assert r["task_is_calls"] == 1 and r["task_is_dims"] == 2 and r["task_is_name"] == "embedding7"
        assert r["weight_calls"] == 1 and r["dims"] == (128, 39_884_406) and r["data_type"] == c["DT_FLOAT"]
        assert r["op_is_this"] and r["initializer_is_ops"] and r["task_points"] == 1 and r["weight_num_dim"] == 2
    assert ref["create_grad"] is True and ref["sync_type"] in (c["PS"], c["NCCL"])
    assert ours["create_grad"] is False and ours["sync_type"] == c["NONE"]
    differ = {k for k in ref if ref[k] != ours[k]}
    assert differ == {"create_grad", "sync_type"}


def test_group_launchers_against_the_references_own_launchers_table_by_table(oracle, rec):
    """The reference's Embedding::forward / backward (src/ops/embedding.cu:274-306, 356-388, object code of
    the reference itself) written down by the same stand-in, one op at a time, against the group launch of
    the same tables: per table the same regions in the same order with the same privileges and field —
    except what the fused mode changes on purpose: no MAP_TO_ZC_MEMORY on output / output-grad, and region 2
    of the backward is the table itself (read-write) where the reference lists the weight-gradient."""
    if not oracle.ref_gpu_available():
        pytest.skip("oracle/_ref/libffref_gpu.so not built")
    GRAD = [0x7E0000000000 + 0x100000000 * t for t in range(4)]
    fwd = rec(0, IDX, OUT, TAB, BAGS, D, B, ROWS)
    bwd = rec(2, IDX, OUT, TAB, BAGS, D, B, ROWS)
    c = fwd["consts"]
    key = lambda q: (q["privilege"], q["base"], q["extent"], q["fields"], q["coherence"])
    for t in range(4):
        one = lambda what: oracle.launch_record("reference", what, IDX[t:t + 1], OUT[t:t + 1], TAB[t:t + 1],
                                                BAGS[t:t + 1], D, B, ROWS[t:t + 1], Wgrad_ptrs=GRAD[t:t + 1])
        rf, rb = one(0), one(2)
        assert rf["task_id"] == c["EMBED_FWD_TASK_ID"] and rb["task_id"] == c["EMBED_BWD_TASK_ID"]
        for r in (rf, rb):  # one point, the op's OpMeta* per point, no by-value argument
            assert r["points"] == 1 and r["local_arg_is_lead_meta"] and r["args"]["len"] == 0 and len(r["reqs"]) == 3
        # forward: identical requirements, but the reference maps the output to zero-copy host memory
        assert [key(q) for q in rf["reqs"]] == [key(q) for q in fwd["reqs"][3 * t:3 * t + 3]]
        assert [q["tag"] for q in rf["reqs"]] == [0, c["MAP_TO_ZC_MEMORY"], 0]
        assert [q["tag"] for q in fwd["reqs"][3 * t:3 * t + 3]] == [0, 0, 0]
        # backward: regions 0 and 1 identical (tag apart); region 2: weight-grad RW there, the table RW here
        ours = bwd["reqs"][3 * t:3 * t + 3]
        assert [key(q) for q in rb["reqs"][:2]] == [key(q) for q in ours[:2]]
        assert [q["tag"] for q in rb["reqs"]] == [0, c["MAP_TO_ZC_MEMORY"], 0]
        assert rb["reqs"][2]["privilege"] == ours[2]["privilege"] == c["READ_WRITE"]
        assert rb["reqs"][2]["extent"] == ours[2]["extent"] == (ROWS[t], D)
        assert rb["reqs"][2]["base"] == GRAD[t] and ours[2]["base"] == TAB[t]
        # the mapping tag FFMapper::slice_task places the task by (src/mapper/mapper.cc:346-438): the group
        # launch carries its LEAD op's — the hash of "embedding0", which is also the name the one-op
        # reference launches above run under — i.e. the device of every op of the group
        assert rf["tag"] == rb["tag"] != 0
        assert fwd["tag"] == bwd["tag"] == rf["tag"]
