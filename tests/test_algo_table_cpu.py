"""Per-bucket all-reduce variant selection from the measured crossover table (no GPU needed)."""
import json

from pytorch_distributed_rnn_b200.parallel import symm


def test_table_loader_reads_world_rows_and_falls_back(tmp_path):
    p = tmp_path / "t.json"
    p.write_text(json.dumps({"worlds": {"8": {"ll_max_bytes": 65536, "oneshot_max_bytes": 1 << 20, "nvls_min_bytes": 1 << 40}}}))
    t8 = symm.load_algo_table(8, str(p))
    assert t8["ll_max_bytes"] == 65536 and t8["oneshot_max_bytes"] == 1 << 20 and t8["source"] == str(p)
    t3 = symm.load_algo_table(3, str(p))          # world size without a measured row: defaults
    assert t3["source"] == "defaults" and t3["ll_max_bytes"] == symm.DEFAULT_TABLE["ll_max_bytes"]
    assert symm.load_algo_table(8, str(tmp_path / "missing.json"))["source"] == "defaults"


def test_committed_table_has_the_measured_worlds():
    t = symm.load_algo_table(8)
    assert t["source"].endswith("profiles/allreduce_table.json"), t
    for w in (2, 4, 8):
        row = symm.load_algo_table(w)
        assert row["ll_max_bytes"] >= 56600      # the reference model's gradient goes through the LL form everywhere


class _Ctx:
    ll_floats = 256 << 10


def _pick(world, table, n, in_heap=True, nvls_ok=False, forced=0):
    c = symm.SymmComm.__new__(symm.SymmComm)       # selection logic only: no device, no region
    c.world, c.table, c.algo, c.ctx = world, table, forced, _Ctx()
    return c.pick_algo(n, in_heap, nvls_ok)


def test_pick_algo_follows_the_table():
    t = {"ll_max_bytes": 65536, "oneshot_max_bytes": 1 << 20, "nvls_min_bytes": 1 << 40}
    assert _pick(8, t, 14150) == symm.ALGO_LL                           # 56.6 KB
    assert _pick(8, t, 100_000) == symm.ALGO_ONESHOT                    # 400 KB: above LL, below the one-shot limit
    assert _pick(8, t, 8_396_800) == symm.ALGO_TWOSHOT                  # a 33.6 MB layer bucket
    assert _pick(8, t, 8_396_800, in_heap=False) == symm.ALGO_ONESHOT   # two-shot needs the symmetric heap
    t2 = dict(t, nvls_min_bytes=1 << 20)
    assert _pick(8, t2, 8_396_800, nvls_ok=True) == symm.ALGO_NVLS
    assert _pick(8, t2, 8_396_800, nvls_ok=False) == symm.ALGO_TWOSHOT
    assert _pick(1, t, 10_000_000) == symm.ALGO_LL or _pick(1, t, 10_000_000) in (symm.ALGO_ONESHOT, symm.ALGO_TWOSHOT)
    assert _pick(8, t, 14150, forced=symm.ALGO_TWOSHOT) == symm.ALGO_TWOSHOT
