"""The size -> variant decision (parallel/comm.py: _VariantTable) on the table that is installed, without a GPU:
what `auto` resolves to for the sizes the examples use, the fall-backs of a forced variant, and a damaged table."""
import json
import os
import warnings

import pytest

from torch_distlearn_b200 import _lib
from torch_distlearn_b200.parallel.comm import _VariantTable

TABLE = os.path.join(os.path.dirname(_lib.__file__), "parallel", "variant_table.json")
FIELDS = ("oneshot_max", "nvls_min", "ll_max", "hybrid_min", "hybrid_permille")
V = _lib.VARIANT_NAMES


@pytest.fixture(autouse=True)
def clean_env(monkeypatch):
    for k in ("DL_B200_VARIANT", "DL_B200_VARIANT_TABLE", "DL_B200_NVLS_PERMILLE", "DL_B200_LIGHT_PERMILLE"):
        monkeypatch.delenv(k, raising=False)


def test_installed_table_has_every_measured_world_and_field():
    data = json.load(open(TABLE))
    for world in ("2", "4", "8"):
        assert world in data, f"world {world} was measured this round"
        for f in FIELDS:
            assert f in data[world], (world, f)
            for op, v in data[world][f].items():
                assert op in _VariantTable.OPS and isinstance(v, int) and v >= 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_auto_on_the_installed_table(world):
    t = _VariantTable(world)
    assert t.source != "defaults"
    cap = 64 * 1024
    # the 10 x 10 confusion matrix (examples/cifar10.lua:203) and anything of 1 KiB goes down the push path
    assert t.pick("allreduce", 800, world, True, "auto", cap) == V["ll"]
    assert t.pick("fused_sgd", 1024, world, True, "auto", cap) == V["ll"]
    # ... unless the communicator could not afford a staging area
    assert t.pick("fused_sgd", 1024, world, True, "auto", 0) != V["ll"]
    # the flagship model's gradient (16.5 MiB): never the one-shot or the push kernel
    big = t.pick("fused_sgd", 4328992 * 4, world, True, "auto", cap, aligned32=True)
    assert big in (V["twoshot"], V["nvls"], V["hybrid"])
    # without a multicast mapping nothing may resolve to the switch-only kernel
    for nbytes in (1 << 10, 1 << 16, 1 << 20, 1 << 24, 1 << 28):
        for op in _VariantTable.OPS:
            assert t.pick(op, nbytes, world, False, "auto", cap, aligned32=True) != V["nvls"]
    assert t.permille("fused_sgd", False) == 0
    assert 0 <= t.permille("fused_sgd", True) <= 1000


def test_two_gpus_keep_the_p2p_kernels_and_the_light_kernel_off_the_switch():
    t = _VariantTable(2)
    assert t.pick("fused_sgd", 1 << 24, 2, True, "auto", 64 * 1024, aligned32=True) == V["twoshot"]
    assert t.light_permille == 0
    assert _VariantTable(8).light_permille == 1000


def test_world_one_never_takes_a_multi_rank_path():
    t = _VariantTable(1)
    for nbytes in (64, 1 << 12, 1 << 20, 1 << 26):
        assert t.pick("fused_sgd", nbytes, 1, False, "auto", 64 * 1024, aligned32=True) in (V["oneshot"], V["twoshot"])


def test_forced_variants_fall_back_where_they_do_not_apply():
    t = _VariantTable(8)
    cap = 64 * 1024
    assert t.pick("fused_sgd", 1 << 20, 8, True, "ll", cap) == V["twoshot"]            # too large for the staging area
    assert t.pick("fused_sgd", 1 << 20, 8, False, "nvls", cap) == V["twoshot"]         # no multicast
    assert t.pick("fused_ea", 1 << 20, 8, True, "oneshot", cap) == V["twoshot"]        # EA has no one-shot form
    assert t.pick("broadcast", 1 << 20, 8, True, "hybrid", cap, aligned32=True) == V["nvls"]   # broadcast has no hybrid form
    assert t.pick("fused_sgd", 1 << 20, 8, True, "hybrid", cap, aligned32=False) == V["nvls"]  # 32-byte items need alignment
    assert t.pick("fused_sgd", 1 << 20, 2, True, "hybrid", cap, aligned32=False) == V["twoshot"]
    assert t.pick("allreduce", 1 << 20, 8, True, "light", cap, aligned32=True) == V["twoshot"]  # light is a fused_sgd kernel
    assert t.pick("fused_sgd", 1 << 20, 8, True, "light", cap, aligned32=True) == V["light"]


def test_environment_overrides(monkeypatch):
    monkeypatch.setenv("DL_B200_VARIANT", "twoshot")
    monkeypatch.setenv("DL_B200_NVLS_PERMILLE", "250")
    monkeypatch.setenv("DL_B200_LIGHT_PERMILLE", "500")
    t = _VariantTable(8)
    assert t.pick("fused_sgd", 1024, 8, True, "auto", 64 * 1024) == V["twoshot"]
    assert t.permille("allreduce", True) == 250 and t.light_permille == 500


def test_a_damaged_table_is_reported_and_the_defaults_stand(tmp_path, monkeypatch):
    bad = tmp_path / "table.json"
    bad.write_text('{"8": {"oneshot_max": {"fused_sgd": "not a number"}}')
    monkeypatch.setenv("DL_B200_VARIANT_TABLE", str(bad))
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        t = _VariantTable(8)
    assert any("variant table" in str(x.message) for x in w)
    assert t.source == "defaults" and t.oneshot_max["fused_sgd"] == _VariantTable.DEFAULT["fused_sgd"]
