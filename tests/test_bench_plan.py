"""Host logic of bench.py that needs no GPU: the per-token mat-vec plan of Llama-3-8B Q4_K_M (tensor-type rules of the reference's
llama-quantize, src/llama-quant.cpp:185-187) and the grouping of its Q4_K mat-vecs into the launches the plugin issues."""
import bench


def test_step_plan_bytes_and_types():
    plan = bench.step_plan()
    assert len(plan) == 32 * 7 + 1
    q4k = [p for p in plan if p[1] == bench.Q4_K]
    q6k = [p for p in plan if p[1] == bench.Q6_K]
    assert len(q4k) == 192 and len(q6k) == 33            # half of attn_v / ffn_down in Q6_K (use_more_bits) + the output matrix
    total = bench.plan_bytes(plan)
    assert abs(total - 4.616e9) < 0.02e9                 # 4.616 GB of weights per decoded token
    assert bench.plan_bytes(plan, only=bench.Q4_K) == 192 * 17498112


def test_issue_groups_cover_every_matrix_once():
    q4k = [p for p in bench.step_plan() if p[1] == bench.Q4_K]
    groups = bench.issue_groups(q4k)
    flat = [e for g in groups for e in g]
    assert flat == q4k                                   # same matrices, same order
    assert len(groups) == 112
    for g in groups:
        assert 1 <= len(g) <= 3
        assert len({e[3] for e in g}) == 1 and len({e[4] for e in g}) == 1      # one K, one activation vector per launch
        names = [e[0] for e in g]
        assert names in (["wq", "wk"], ["wq", "wk", "wv"], ["wo"], ["gate", "up"], ["down"]), names
