"""The pass schedule (identical on every rank) has to add up: generations, halo depths, and
every halo push must be consumed inside the same step() call (CPU only, no device)."""
import pytest
from hypothesis import given, settings, strategies as st

from gol_b200 import GolError
from gol_b200._native import describe_schedule


def check(sched, gens):
    items = sched["items"]
    if gens == 0:
        assert items == []
        return
    assert items[0]["prime"] == 1
    pending = items[0]["prime_depth"]      # rows waiting in the neighbours' recv buffers
    ext = 0                                # halo rows currently valid around the interior
    total = 0
    for it in items[1:]:
        assert it["prime"] == 0
        if it["consume"]:
            assert pending == it["ext_in"] > 0 and ext == 0
            ext, pending = it["ext_in"], 0
        else:
            assert it["ext_in"] == ext
        assert 1 <= it["k"] <= ext
        ext -= it["k"]
        total += it["k"]
        if it["push"]:
            assert ext == 0 and pending == 0     # only the last pass of a period pushes
            pending = it["push"]
    assert total == gens and pending == 0 and ext == 0


@settings(max_examples=200, deadline=None)
@given(gens=st.integers(0, 300), k=st.sampled_from([1, 2, 3, 4, 6, 8]), mult=st.sampled_from([1, 2, 4]),
       rows=st.integers(32, 500))
def test_fused_schedule_adds_up(gens, k, mult, rows):
    depth = k * mult
    if depth > 32:
        depth = k
    s = describe_schedule(1024, rows, gens, mode="fused", k=k, depth=depth)
    assert s["plan"]["temporal_k"] == k and s["plan"]["exchange_depth"] == depth
    check(s, gens)
    assert all(it["k"] in (1, 2, 3, 4, 6, 8) for it in s["items"][1:])


def test_generic_schedule_is_one_generation_per_pass():
    s = describe_schedule(100, 7, 5, mode="auto")
    assert s["plan"]["mode"] == 1 and len(s["items"]) == 6
    check(s, 5)
    assert [it["push"] for it in s["items"][1:]] == [1, 1, 1, 1, 0]


def test_auto_shrinks_depth_for_short_slabs_and_rejects_bad_requests():
    s = describe_schedule(128, 3, 10, mode="auto")
    assert s["plan"]["mode"] == 2 and s["plan"]["exchange_depth"] <= 3
    check(s, 10)
    with pytest.raises(GolError, match="multiple of 64"):
        describe_schedule(100, 64, 4, mode="fused")
    with pytest.raises(GolError, match="multiple of temporal_k"):
        describe_schedule(128, 64, 4, mode="fused", k=4, depth=6)
    with pytest.raises(GolError, match="exceeds the slab height"):
        describe_schedule(128, 8, 4, mode="fused", k=4, depth=16)
