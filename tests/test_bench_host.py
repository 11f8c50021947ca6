"""Host-side logic of bench.py that needs no GPU."""
import ast
import os

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sustained_plan_depends_on_the_hint_only():
    """Two ranks with slightly different step times must not plan different chunk counts: the
    plan is computed from ONE hint (bench.py reduces it over the ranks first)."""
    a = bench.sustained_plan(3.0, 0.1788)
    b = bench.sustained_plan(3.0, 0.1788)
    assert a == b and a[0] >= 1 and a[1] >= 3
    # what went wrong at N = 8 (profiles/r02p_scale_n8_failure_excerpt.txt): per-rank hints a
    # few percent apart give different chunk counts
    plans = {bench.sustained_plan(3.0, ms)[1] for ms in (0.170, 0.175, 0.180, 0.185, 0.190)}
    assert len(plans) > 1
    chunk, n = bench.sustained_plan(3.0, 1.3)
    assert 2.5e3 <= chunk * n * 1.3 <= 4.0e3          # about three seconds of work


def test_the_hint_is_reduced_over_ranks_before_the_plan():
    """run_ours must all-reduce the hint (MAX) before calling sustained_leg."""
    src = open(os.path.join(ROOT, "bench.py")).read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "run_ours")
    body = ast.get_source_segment(src, fn)
    i_reduce = body.index("dist.all_reduce(hint")
    i_leg = body.index("sustained_leg(")
    assert i_reduce < i_leg
    assert "float(hint[0])" in body[i_leg:i_leg + 400]


def test_kv_bytes_of_the_baseline_configs():
    """sum(seq_len) * Hkv * D * 2 * sizeof(cache_t)."""
    assert bench.kv_bytes(bench.seq_lens_of("cfg4"), 8, "fp8") == 128 * 32768 * 8 * 128 * 2
    assert bench.kv_bytes(bench.seq_lens_of("cfg2"), 8, "auto") == 64 * 4096 * 8 * 128 * 2 * 2
    lens = bench.seq_lens_of("cfg5")
    assert len(lens) == 64 and min(lens) >= 128 and max(lens) <= 131072
