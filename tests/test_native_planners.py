"""Host-side planning of the native library, exercised WITHOUT a GPU: geometry / eligibility queries run no device
work (the SM count falls back to a B200's 148 where no device is present), so their decisions are pinned on CPU."""
import pytest

from tiresias_b200.ops import native


@pytest.fixture(scope="module")
def lib():
    return native.lib()


def test_conv3x3_geometry_queries(lib):
    # ResNet-50 at batch 256 (flat layout), VGG-16 at batch 128 (segment mode from 100 columns on), AlexNet
    for n, hw, cin, cout in ((256, 56, 64, 64), (256, 28, 128, 128), (256, 14, 256, 256), (256, 7, 512, 512),
                             (128, 224, 64, 64), (128, 112, 128, 128), (128, 56, 256, 256), (256, 13, 192, 384)):
        assert native.conv3x3_supported(n, hw, hw, cin, cout), (n, hw, cin, cout)
        assert native.conv3x3_supported(n, hw, hw, cout, cin, dgrad=True)
    assert not native.conv3x3_supported(8, 32, 32, 3, 64)            # 3 input channels: not a multiple of 64
    assert not native.conv3x3_supported(8, 32, 32, 96, 64)
    assert not native.conv3x3_supported(8, 32, 32, 64, 60)           # output channels must be a multiple of 8
    assert not native.conv3x3_supported(0, 32, 32, 64, 64)


def test_halo_wgrad_workspace_follows_the_split_plan(lib):
    ws = lib.tb200_conv3x3_wgrad_workspace
    # items = 2 tap halves x Cin/64 x ceil(Cout/128); splits = 148 // items; one fp32 plane per split
    assert ws(256, 56, 56, 64, 64) == 74 * 64 * 9 * 64 * 4
    assert ws(256, 28, 28, 128, 128) == 37 * 128 * 9 * 128 * 4
    assert ws(256, 14, 14, 256, 256) == 9 * 256 * 9 * 256 * 4
    assert ws(256, 7, 7, 512, 512) == 2 * 512 * 9 * 512 * 4
    assert ws(8, 28, 28, 96, 128) == 0                             # Cin not a multiple of 64
    assert ws(8, 28, 28, 64, 192) == 0                             # Cout neither 64 nor a multiple of 128
    assert ws(128, 224, 224, 64, 64) == 0                          # a stage of whole 225-pixel rows does not fit


def test_gemm_split_plan(lib):
    plan = lib.tb200_gemm_plan_splits
    assert plan(8192, 8192, 8192) == 1                             # the tile grid fills the machine
    assert plan(802816, 256, 64) == 1
    assert plan(64, 4096, 25088) > 1                               # tiny batch, long reduction: split-K
    assert plan(256, 4096, 25088) > 1
    s = plan(64, 1000, 2048)
    assert 1 <= s <= 148
    assert plan(0, 16, 16) == 0


def test_small_kernel_eligibility_queries(lib):
    assert all(lib.tb200_layernorm_supported(d) == 1 for d in (256, 512, 768, 1024))
    assert all(lib.tb200_layernorm_supported(d) == 0 for d in (128, 384, 2048, 500))
    assert lib.tb200_maxpool_supported(3, 2, 1) == lib.tb200_maxpool_supported(2, 2, 0) == lib.tb200_maxpool_supported(3, 2, 0) == 1
    assert lib.tb200_maxpool_supported(3, 1, 1) == 0 and lib.tb200_maxpool_supported(5, 2, 2) == 0
    # the ReLU bit mask needs the streamed BatchNorm kernels: power-of-two channel counts, tensors worth streaming
    assert lib.tb200_bn_mask_supported(802816, 256) == 1
    assert lib.tb200_bn_mask_supported(802816, 72) == 0
    assert lib.tb200_bn_mask_supported(16, 64) == 0
    assert lib.tb200_layernorm_scratch_floats(512) == 148 * 4 * 2 * 512


def test_argument_errors_are_reported_not_crashed(lib):
    """Validation that runs before any device work: every refusal returns non-zero and leaves a readable message
    (two of these messages had a malformed format in round 2 and crashed the caller instead)."""
    cases = [
        (lambda: lib.tb200_conv3x3_fprop(0, 0, 0, 2, 8, 8, 96, 64, 0, 0, None, 0), b"multiple of 64"),
        (lambda: lib.tb200_conv3x3_dgrad(0, 0, 0, 2, 8, 8, 64, 60, 0), b"multiple of 64"),
        (lambda: lib.tb200_conv3x3_wgrad(0, 0, 0, 2, 8, 8, 64, 64, 0, 0, 0), b"workspace too small"),
        (lambda: lib.tb200_conv3x3_wgrad(0, 0, 0, 2, 8, 8, 96, 64, 0, 0, 0), b"multiples of 64"),
        (lambda: lib.tb200_conv3x3_wgrad(0, 0, 0, 2, 8, 8, 64, 192, 0, 0, 0), b"Cout (192)"),
        (lambda: lib.tb200_attention_fwd(0, 0, 0, 0, 128, 8, 1, 0), b"empty problem"),
        (lambda: lib.tb200_attention_fwd(0, 0, 0, 70000, 128, 8, 1, 0), b"too large"),
        (lambda: lib.tb200_layernorm_fwd(0, 0, 0, 0, 0, 0, 0, 16, 500, 1e-5, 0), b"D (500)"),
        (lambda: lib.tb200_layernorm_bwd(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 16, 512, 0), b"scratch too small"),
        (lambda: lib.tb200_maxpool_fwd(0, 0, 0, 2, 8, 8, 8, 5, 2, 2, 0), b"geometry not served"),
        (lambda: lib.tb200_maxpool_fwd(0, 0, 0, 2, 8, 8, 7, 3, 2, 1, 0), b"channels (7)"),
        (lambda: lib.tb200_gemm_bf16_add(0, 0, 0, 0, 16, 16, 16, 0, 0, 0), b"null addend"),
        (lambda: lib.tb200_conv2d_fprop(0, 0, 0, 2, 8, 8, 3, 64, 3, 3, 1, 1, 0), b"channel count (3)"),
    ]
    for call, needle in cases:
        assert call() != 0
        assert needle in lib.tb200_last_error(), (needle, lib.tb200_last_error())
