"""CPU model of the halo-tile convolution's geometry (csrc/conv_halo.cu), step for step: padded-space
tiling, whole-padded-row loads with zero fill outside the image, taps as windows shifted by r*Wp+s rows of
the SAME tile, and stores clipped against the output's real extent. The kernel itself has not run on
hardware yet; this pins its index arithmetic against a plain convolution so that bring-up on the GPU only
has to debug the PTX plumbing."""
import pytest
import torch


def halo_conv_model(x_nhwc: torch.Tensor, w_krsc: torch.Tensor) -> torch.Tensor:
    n, h, w, c = x_nhwc.shape
    k, r_, s_, _ = w_krsc.shape
    pad = (r_ - 1) // 2
    wp = 8
    while wp < w + s_ - 1:
        wp *= 2
    assert wp <= 128
    hp, th = h + r_ - 1, 128 // wp
    a_rows = th + r_ - 1
    num_m = (n * hp + th - 1) // th
    store_w = min(32, wp)
    a_alloc = (a_rows * wp + (s_ - 1) + 7) // 8 * 8
    y = torch.full((n, h, w, k), float("nan"))
    for m_tile in range(num_m):
        acc = torch.zeros(128, k)
        for cb in range(c // 64):
            # producer: a_rows whole padded rows {64 ch, Wp, 1, 1} at (cb*64, -pad, hp-pad, n); the rest of the
            # stage is whatever was there before (poisoned here: only junk outputs may touch it)
            tile = torch.full((a_alloc, 64), 1e30)
            row0 = m_tile * th
            for i in range(a_rows):
                rg = row0 + i
                img, hrow = divmod(rg, hp)
                for j in range(wp):
                    hh, ww = hrow - pad, j - pad
                    inside = img < n and 0 <= hh < h and 0 <= ww < w
                    tile[i * wp + j] = x_nhwc[img, hh, ww, cb * 64:(cb + 1) * 64] if inside else 0.0
            # MMA issuer: per tap a 128-row window starting (r*Wp + s) rows into the tile
            for t in range(r_ * s_):
                r, s = divmod(t, s_)
                win = tile[r * wp + s: r * wp + s + 128]
                acc += win @ w_krsc[:, r, s, cb * 64:(cb + 1) * 64].t()
        # epilogue: per warp 32 rows, TMA stores of store_w positions clipped against (W, H, N)
        for ew in range(4):
            for seg in range(32 // store_w):
                q = m_tile * 128 + ew * 32 + seg * store_w
                rg, j = divmod(q, wp)
                img, hrow = divmod(rg, hp)
                for d in range(store_w):
                    if img < n and hrow < h and j + d < w:
                        y[img, hrow, j + d] = acc[ew * 32 + seg * store_w + d]
    return y


@pytest.mark.parametrize("shape", [(2, 64, 8, 14, 3), (1, 128, 16, 28, 3), (3, 64, 8, 7, 3), (2, 64, 8, 12, 3),
                                   (1, 64, 8, 56, 3), (2, 64, 8, 9, 1)],
                         ids=lambda s: "x".join(map(str, s)))
def test_halo_geometry_equals_convolution(shape):
    n, c, k, hw, r = shape
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, hw, hw, c, generator=g)
    w = torch.randn(k, r, r, c, generator=g) / (c * r * r) ** 0.5
    y = halo_conv_model(x, w)
    ref = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), padding=(r - 1) // 2).permute(0, 2, 3, 1)
    assert not torch.isnan(y).any(), "some real output position was never stored"
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-4), float((y - ref).abs().max())


@pytest.mark.parametrize("k,pad", [(3, 1), (1, 0), (5, 2)])
def test_dgrad_is_a_forward_convolution_with_the_flipped_filter(k, pad):
    """ops.native.conv2d_dgrad_halo runs the forward kernel over dy with dgrad_weight(w): check that identity
    against autograd on the CPU."""
    from tiresias_b200.ops.native import dgrad_weight
    g = torch.Generator().manual_seed(2)
    x = torch.randn(2, 8, 9, 9, generator=g, requires_grad=True)
    w = torch.randn(6, 8, k, k, generator=g)
    y = torch.nn.functional.conv2d(x, w, padding=pad)
    dy = torch.randn(y.shape, generator=g)
    y.backward(dy)
    dx = torch.nn.functional.conv2d(dy, dgrad_weight(w), padding=k - 1 - pad)
    assert torch.allclose(dx, x.grad, rtol=1e-4, atol=1e-4)


def wgrad_flat_model(x_nhwc: torch.Tensor, dy_nhwc: torch.Tensor, splits: int = 3) -> torch.Tensor:
    """csrc/conv3x3.cu conv3x3_wgrad_kernel, step for step: super blocks of 128 flat padded positions (pitch
    W+1 / H+1), a stage = the padded rows of x covering the block plus its halo and the padded rows of dy covering
    the block (rows outside the tensor arrive as zeros, anything else in the stage is poison), a tap = the x
    window shifted by (r-1)*Wp + (s-1) rows, splits over the positions summed in order."""
    n, h, w, cin = x_nhwc.shape
    cout = dy_nhwc.shape[3]
    wp, hp = w + 1, h + 1
    total = n * hp * wp
    nrx = (3 * wp + 129 + wp - 1) // wp
    nry = (wp + 127 + wp - 1) // wp
    num_sb = (total + 127) // 128

    def padded_row(t, rho):                       # one TMA box {C, Wp, 1, 1} at (0, 0, hp, img)
        out = torch.zeros(wp, t.shape[3])
        if rho >= 0:
            img, hrow = divmod(rho, hp)
            if img < n and hrow < h:
                out[:w] = t[img, hrow]
        return out

    planes = []
    for ks in range(splits):
        acc = torch.zeros(9, cout, cin)
        for sb in range(ks * num_sb // splits, (ks + 1) * num_sb // splits):
            q0 = sb * 128
            rho0 = (q0 + wp - 1) // wp - 2
            rhoy = q0 // wp
            xs = torch.full((nrx * wp + 8, cin), 1e30)
            ys = torch.full((nry * wp + 8, cout), 1e30)
            for i in range(nrx):
                xs[i * wp:(i + 1) * wp] = padded_row(x_nhwc, rho0 + i)
            for i in range(nry):
                ys[i * wp:(i + 1) * wp] = padded_row(dy_nhwc, rhoy + i)
            a0 = q0 - rhoy * wp
            b_mid = q0 - rho0 * wp
            assert 0 <= a0 and a0 + 128 <= nry * wp
            for t in range(9):
                r, s = divmod(t, 3)
                b0 = b_mid + (r - 1) * wp + (s - 1)
                assert 0 <= b0 and b0 + 128 <= nrx * wp
                acc[t] += ys[a0:a0 + 128].t() @ xs[b0:b0 + 128]
        planes.append(acc)
    dw = sum(planes)                              # [9, cout, cin]
    return dw.permute(1, 0, 2).reshape(cout, 3, 3, cin)


@pytest.mark.parametrize("n,h,w", [(2, 7, 7), (1, 14, 14), (3, 5, 9), (1, 28, 28), (2, 3, 56), (1, 1, 1), (1, 2, 130)])
def test_wgrad_flat_space_geometry_equals_weight_gradient(n, h, w):
    g = torch.Generator().manual_seed(3)
    cin, cout = 4, 5
    x = torch.randn(n, h, w, cin, generator=g)
    dy = torch.randn(n, h, w, cout, generator=g)
    got = wgrad_flat_model(x, dy)
    ref = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2), (cout, cin, 3, 3), dy.permute(0, 3, 1, 2), stride=1, padding=1)
    assert torch.allclose(got, ref.permute(0, 2, 3, 1), rtol=1e-4, atol=1e-4), (got - ref.permute(0, 2, 3, 1)).abs().max()
