"""Pure-Python pieces of the tensor-core convolution path (no GPU): geometry gate and split-K choice."""
from torch_distlearn_b200.ops.conv import _implicit_ok, _splits


def test_implicit_gemm_geometry_gate():
    for (B, H, cin, cout) in [(128, 16, 64, 128), (128, 8, 128, 256), (128, 4, 256, 512), (8, 4, 256, 512), (1, 16, 32, 32)]:
        assert _implicit_ok(B, H, H, cin, cout), (B, H, cin, cout)
    assert not _implicit_ok(128, 32, 32, 3, 64)          # Cin = 3: explicit im2col path
    assert not _implicit_ok(4, 12, 12, 64, 64)           # 12-pixel rows do not tile the 32 / 128 pixel TMA boxes
    assert not _implicit_ok(1, 4, 4, 64, 64)             # 16 pixels: not a whole 32-pixel K block for wgrad
    assert not _implicit_ok(2, 16, 16, 48, 64)           # channels must be multiples of 32


def test_split_k_fills_the_machine_without_starving_splits():
    assert _splits(256, 1, 1600) == 1                    # layer-2 forward: 256 tiles already cover 148 SMs
    assert _splits(1024, 1, 96) == 1
    s = _splits(1, 13, 32768)                            # layer-2 wgrad: 13 tiles, K = 32768 -> about one wave of CTAs
    assert 8 <= s <= 12 and 13 * s <= 148
    for tiles, k in [(1, 128), (3, 64), (50, 8192), (200, 2048)]:
        s = _splits(tiles, 1, k)
        assert s >= 1 and (k // 32) // s >= 4 or s == 1  # every split keeps >= 4 K blocks
