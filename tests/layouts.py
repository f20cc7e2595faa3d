"""Test helpers: encode / decode the spectral-internal device layouts documented in include/makani_b200.h
with plain torch ops (independent of the CUDA layout kernels)."""
import torch


def rup(x, m):
    return (x + m - 1) // m * m


def xm_decode(xm, B, C, K, M):
    """XM flat [M][B][2][Cp][Kp] -> complex [B, C, K, M]."""
    Cp, Kp = rup(C, 4), rup(K, 4)
    v = xm.view(M, B, 2, Cp, Kp)[:, :, :, :C, :K]
    return torch.complex(v[:, :, 0], v[:, :, 1]).permute(1, 2, 3, 0).contiguous()


def xm_encode(c, garbage=float("nan")):
    """complex [B, C, K, M] -> XM flat; pad entries are filled with `garbage` (they must never matter)."""
    B, C, K, M = c.shape
    Cp, Kp = rup(C, 4), rup(K, 4)
    out = torch.full((M, B, 2, Cp, Kp), garbage, dtype=torch.float32, device=c.device)
    out[:, :, 0, :C, :K] = c.real.permute(3, 0, 1, 2)
    out[:, :, 1, :C, :K] = c.imag.permute(3, 0, 1, 2)
    # pad latitudes of real channels must be finite zeros (TMA never reads them, but keep the contract)
    return out.reshape(-1)


def sp_decode(sp, B, C, L, M):
    """SP flat [L][M][B][2][Cp] -> complex [B, C, L, M]; entries with m > l are forced to zero (undefined region)."""
    Cp = rup(C, 4)
    v = sp.view(L, M, B, 2, Cp)[..., :C]
    c = torch.complex(v[..., 0, :], v[..., 1, :]).permute(2, 3, 0, 1).contiguous()
    l = torch.arange(L, device=sp.device).view(L, 1)
    m = torch.arange(M, device=sp.device).view(1, M)
    return torch.where(m <= l, c, torch.zeros_like(c))


def sp_encode(c, garbage=float("nan")):
    """complex [B, C, L, M] -> SP flat.  Only what the layout contract defines is filled in: real channels, for
    m < min(M, round_up(l+1,16)), with exact zeros where m > l.  Everything else (pad channels, the undefined
    region) is `garbage`, so a kernel that lets it leak into a result shows up as NaN."""
    B, C, L, M = c.shape
    Cp = rup(C, 4)
    l = torch.arange(L, device=c.device).view(L, 1)
    m = torch.arange(M, device=c.device).view(1, M)
    c = torch.where(m <= l, c, torch.zeros_like(c))
    out = torch.full((L, M, B, 2, Cp), garbage, dtype=torch.float32, device=c.device)
    out[..., 0, :C] = c.real.permute(2, 3, 0, 1)
    out[..., 1, :C] = c.imag.permute(2, 3, 0, 1)
    defined = m < torch.clamp(((l + 16) // 16) * 16, max=M)
    und = (~defined).view(L, M, 1, 1, 1).expand_as(out)
    out = torch.where(und, torch.full_like(out, garbage), out)
    return out.reshape(-1)


def wp_decode(wp, Cin, Cout, L):
    """WP flat [L][2][Coutp][Cinp] -> complex [Cin, Cout, L]."""
    Cinp, Coutp = rup(Cin, 4), rup(Cout, 4)
    v = wp.view(L, 2, Coutp, Cinp)[:, :, :Cout, :Cin]
    return torch.complex(v[:, 0], v[:, 1]).permute(2, 1, 0).contiguous()


def tri_mask(L, M, device):
    l = torch.arange(L, device=device).view(L, 1)
    m = torch.arange(M, device=device).view(1, M)
    return (m <= l)


def rel_err(a, b):
    a = a.double() if not a.is_complex() else a.to(torch.complex128)
    b = b.to(a.dtype)
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def tf32_rn(t):
    """Round-to-nearest (ties away) to TF32, bit-exact with cvt.rna.tf32.f32 for finite values."""
    if t.is_complex():
        return torch.complex(tf32_rn(t.real.contiguous()), tf32_rn(t.imag.contiguous()))
    u = t.contiguous().view(torch.int32)
    return ((u + 0x1000) & ~0x1FFF).view(torch.float32)
