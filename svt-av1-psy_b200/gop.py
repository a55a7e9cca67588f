"""Synthetic random-access prediction structure for the ME benchmark.

Picture Decision (pd_process.c) is the producer of the ME stage's inputs and is out of scope; the
benchmark still needs plausible reference lists, so this module lays out a hierarchical-B mini-GOP
the way the reference's random-access structures do in outline: pictures of temporal layer t
reference the nearest already-available pictures of lower layers, list 0 in the past, list 1 in the
future; base-layer pictures reference earlier base pictures in both lists (for them the reference
skips HME for list 1 and mirrors list 0, motion_estimation.c:1839-1848). Reference counts follow
scs->mrp_ctrls (enc_handle.c:3799-3960)."""


def mini_gop_size(hier_levels):
    return 1 << hier_levels


def temporal_layer(k, hier_levels):
    """k = 1..mini_gop_size: position inside the mini-GOP (k == size is the base picture)"""
    size = 1 << hier_levels
    if k == size:
        return 0
    tz = (k & -k).bit_length() - 1
    return hier_levels - tz


def picture_plan(poc, hier_levels, max_l0, max_l1, n_frames):
    """-> dict(poc, tl, is_ref, l0=[pocs], l1=[pocs]) for poc >= 1 (poc 0 is the key frame: no ME)"""
    size = 1 << hier_levels
    g, k = divmod(poc - 1, size)
    k += 1
    start = g * size  # previous base (or key frame)
    tl = temporal_layer(k, hier_levels)
    if tl == 0:
        past = [p for p in range(start, -1, -size)][:max_l0]
        return dict(poc=poc, tl=0, is_ref=1, l0=past, l1=past[:max_l1])
    step = 1 << (hier_levels - tl)  # distance to the neighbours of the next lower layer
    avail = [start + j for j in range(0, size + 1) if j == 0 or j == size or temporal_layer(j, hier_levels) < tl]
    avail = [p for p in avail if p < n_frames]
    older = [p for p in range(start - size, -1, -size)][:2]  # earlier bases stay in the DPB
    past = sorted([p for p in avail if p < poc] + older, key=lambda p: poc - p)[:max_l0]
    future = sorted([p for p in avail if p > poc], key=lambda p: p - poc)[:max_l1]
    if not future:  # tail of the clip: no future picture available -> P-like, list 1 = list 0
        future = past[:max_l1]
    del step
    return dict(poc=poc, tl=tl, is_ref=1 if tl < hier_levels else 0, l0=past, l1=future)


def coding_order(first_poc, hier_levels, n_frames):
    """pictures of the mini-GOP starting after base `first_poc`, in an order in which every reference is
    available: base first, then by temporal layer"""
    size = 1 << hier_levels
    pocs = [p for p in range(first_poc + 1, first_poc + size + 1) if p < n_frames]
    return sorted(pocs, key=lambda p: (temporal_layer(p - first_poc, hier_levels), p))


def mrp_for_preset(enc_mode):
    """The scs->mrp_ctrls fields the ME stage reads, for CRF random access, non-screen content, above
    360p, fast_decode 0 (mrp level table enc_handle.c:3799-3960, level selection :4588-4613)."""
    if enc_mode <= -1:  # MR and slower: mrp_level 1
        return dict(only_l_bwd=0, safe_limit_nref=0, safe_limit_zz_th=0, l0=4, l1=3)
    if enc_mode <= 4:  # mrp_level 2 (M0-M1) / 5 (M2-M4)
        return dict(only_l_bwd=1, safe_limit_nref=0, safe_limit_zz_th=0, l0=4, l1=3)
    if enc_mode <= 8:  # mrp_level 9
        return dict(only_l_bwd=1, safe_limit_nref=2, safe_limit_zz_th=0, l0=3, l1=2)
    return dict(only_l_bwd=1, safe_limit_nref=2, safe_limit_zz_th=0, l0=2, l1=2)  # mrp_level 10
