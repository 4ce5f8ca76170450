"""The host-side planner of libeasyav1_cuda (csrc/sched.h through eav1cu_plan_units): which
commands become independent motion-compensation items, which become dataflow units, which
consecutive commands fuse, and which units wait for which. Needs no GPU."""
import ctypes as C
import struct

from easyav1_b200.lib import load_engine

OP = dict(NOP=0, MC=1, WARP=2, AVG=3, W_AVG=4, MASK=5, W_MASK=6, BLEND=7, BLEND_V=8, BLEND_H=9, IPRED=10,
          CFL=11, PAL=12, ITX=13)
F_ARENA, F_PREP, F_HEAD, F_MASK_ARENA = 1, 2, 4, 8


def cmd(op, plane, w, h, x=0, y=0, dst_off=0, flags=0, u=b""):
    """mirror of struct Eav1Cmd (include/easyav1_cuda.h): 16-byte head + 44-byte union"""
    head = struct.pack("<BBBBhhIHBB", OP[op], plane, w, h, x, y, dst_off, w, flags, 0)
    assert len(head) == 16
    return head + u.ljust(44, b"\0")


def u_mc(ref=0, sx=0, sy=0):
    return struct.pack("<iiiiHHHHBBBB", sx, sy, 1 << 24, 1 << 24, 0, 0, 0, 0, ref, 0, 0, 0)


def u_ipred(ex, ey, have=3, n_left=8, n_top=8, tw=2, th=2):
    return struct.pack("<hhhBBBBBBBBhh", ex, ey, 0, 0, have, n_left, 0, n_top, 0, tw, th, 64, 64)


def u_comp(t1, t2, mask_off=0):
    return struct.pack("<IIIhBBB", t1, t2, mask_off, 0, 0, 1, 1)


def plan(cmds, w=64, h=64):
    E = load_engine()
    E.eav1cu_plan_units.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                    C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.c_size_t,
                                    C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_size_t)]
    buf = b"".join(cmds)
    assert len(buf) == 60 * len(cmds)
    pw = (C.c_int * 3)(w, w // 2, w // 2)
    ph = (C.c_int * 3)(h, h // 2, h // 2)
    cap = 64
    first, cnt, lvl, items = ((C.c_uint32 * cap)() for _ in range(4))
    n_items = C.c_size_t(0)
    n = E.eav1cu_plan_units(buf, len(cmds), pw, ph, first, cnt, lvl, cap, items, cap, C.byref(n_items))
    assert n >= 0
    units = [(first[i], cnt[i], lvl[i]) for i in range(n)]
    return units, [items[i] for i in range(n_items.value)]


def test_struct_size():
    E = load_engine()
    E.eav1cu_abi_sizeof.restype = C.c_size_t
    assert E.eav1cu_abi_sizeof(0) == len(cmd("NOP", 0, 4, 4)) == 60


def test_reference_mc_is_phase_one_and_tiled():
    units, items = plan([cmd("MC", 0, 64, 64, 0, 0, flags=F_HEAD, u=u_mc()),
                         cmd("MC", 1, 32, 32, 0, 0, u=u_mc())])
    assert units == []
    # 64x64 -> four 32x32 tiles (tile_x << 24 | tile_y << 26), 32x32 -> one item
    assert sorted(items) == sorted([0, 1 << 24, 1 << 26, (1 << 24) | (1 << 26), 1])


def test_intrabc_mc_is_a_dataflow_unit():
    units, items = plan([cmd("IPRED", 0, 8, 8, 0, 0, flags=F_HEAD, u=u_ipred(0, 0, have=0)),
                         cmd("MC", 0, 8, 8, 16, 0, flags=F_HEAD, u=u_mc(ref=7, sx=0, sy=0))])
    assert items == []
    assert units == [(0, 2, 0)]                      # it consumes the block before it: same unit, in order


def test_prediction_and_residual_fuse():
    units, _ = plan([cmd("IPRED", 0, 8, 8, 0, 0, flags=F_HEAD, u=u_ipred(0, 0, have=0)),
                     cmd("ITX", 0, 8, 8, 0, 0),
                     cmd("IPRED", 0, 8, 8, 8, 0, u=u_ipred(8, 0, have=1)),      # left edge = first block
                     cmd("ITX", 0, 8, 8, 8, 0)])
    assert units == [(0, 4, 0)]                      # one unit, four commands run back to back


def test_neighbour_blocks_form_a_wavefront():
    c = []
    for i in range(4):                               # four blocks in a row, each needs its left neighbour
        c.append(cmd("IPRED", 0, 8, 8, 8 * i, 0, flags=F_HEAD, u=u_ipred(8 * i, 0, have=1 if i else 0, n_top=0)))
        c.append(cmd("ITX", 0, 8, 8, 8 * i, 0))
    c.append(cmd("IPRED", 1, 4, 4, 0, 0, flags=F_HEAD, u=u_ipred(0, 0, have=0)))   # chroma: independent
    units, _ = plan(c)
    # block i+1 consumes block i directly, so the whole row fuses; the chroma block stands alone at level 0
    assert units[0][1] == 8 and units[0][2] == 0
    assert units[-1] == (8, 1, 0)


def test_independent_residuals_stay_parallel():
    units, items = plan([cmd("MC", 0, 16, 16, 0, 0, flags=F_HEAD, u=u_mc()),
                         cmd("ITX", 0, 8, 8, 0, 0), cmd("ITX", 0, 8, 8, 8, 0), cmd("ITX", 0, 8, 8, 0, 8)])
    assert items == [0]
    assert units == [(1, 1, 0), (2, 1, 0), (3, 1, 0)]


def test_mask_waits_for_w_mask_through_the_arena():
    units, items = plan([cmd("MC", 0, 16, 16, dst_off=0, flags=F_HEAD | F_ARENA | F_PREP, u=u_mc()),
                         cmd("MC", 0, 16, 16, dst_off=512, flags=F_ARENA | F_PREP, u=u_mc(ref=1)),
                         cmd("W_MASK", 0, 16, 16, 0, 0, u=u_comp(0, 512, mask_off=4096)),
                         cmd("MC", 1, 8, 8, dst_off=1024, flags=F_ARENA | F_PREP, u=u_mc()),
                         cmd("MC", 1, 8, 8, dst_off=1280, flags=F_ARENA | F_PREP, u=u_mc(ref=1)),
                         cmd("MASK", 1, 8, 8, 0, 0, flags=F_MASK_ARENA, u=u_comp(1024, 1280, mask_off=4096))])
    assert sorted(items) == [0, 1, 3, 4]
    assert units == [(2, 1, 0), (5, 1, 1)]           # chroma mask blend is one level after w_mask


def test_mc_over_already_written_cells_is_not_hoisted():
    units, items = plan([cmd("IPRED", 0, 8, 8, 0, 0, flags=F_HEAD, u=u_ipred(0, 0, have=0)),
                         cmd("MC", 0, 8, 8, 0, 0, flags=F_HEAD, u=u_mc())])
    assert items == []
    assert len(units) in (1, 2) and sum(u[1] for u in units) == 2


def tickets(cmds, w=64, h=64, est_weight=8):
    E = load_engine()
    E.eav1cu_plan_tickets.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int,
                                      C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_uint32),
                                      C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_size_t)]
    buf = b"".join(cmds)
    pw = (C.c_int * 3)(w, w // 2, w // 2)
    ph = (C.c_int * 3)(h, h // 2, h // 2)
    cap = 4096
    order, dep_start, deps = (C.c_uint32 * cap)(), (C.c_uint32 * (cap + 1))(), (C.c_uint32 * (8 * cap))()
    nd = C.c_size_t(0)
    n = E.eav1cu_plan_tickets(buf, len(cmds), pw, ph, est_weight, order, cap, dep_start, deps, 8 * cap, C.byref(nd))
    assert n >= 0
    return [order[i] for i in range(n)], [[deps[k] for k in range(dep_start[u], dep_start[u + 1])] for u in range(n)]


def wavefront_frame(rows=6, cols=8, chroma_free=5):
    """rows x cols intra blocks, each needing its left, top and top-right neighbours (two separate
    commands per block so nothing fuses across blocks), plus independent chroma blocks"""
    c = []
    for r in range(rows):
        for k in range(cols):
            have = (1 if k else 0) | (2 if r else 0) | (16 if r and k + 1 < cols else 0)
            c.append(cmd("IPRED", 0, 8, 8, 8 * k, 8 * r, flags=F_HEAD,
                         u=u_ipred(8 * k, 8 * r, have=have, n_left=8 if k else 0, n_top=8 if r else 0)))
            c.append(cmd("ITX", 0, 8, 8, 8 * k, 8 * r))
    for i in range(chroma_free):
        c.append(cmd("IPRED", 1, 4, 4, 4 * i, 0, flags=F_HEAD, u=u_ipred(4 * i, 0, have=0)))
    return c


def test_ticket_order_is_a_permutation_and_producers_come_first():
    """The progress argument of recon_frame_kernel: a unit only waits for smaller tickets."""
    for weight in (1, 2, 8):
        order, deps = tickets(wavefront_frame(), est_weight=weight)
        assert sorted(order) == list(range(len(order)))
        ticket_of = {u: t for t, u in enumerate(order)}
        assert any(deps), "the fixture must contain dependent units"
        for u, prods in enumerate(deps):
            for p in prods:
                assert p < u                               # decode order is topological
                assert ticket_of[p] < ticket_of[u]         # and so is the ticket order


def test_ticket_order_leads_with_the_critical_chain():
    """Independent units must not be queued in front of the second unit of a long chain."""
    c = []
    for i in range(40):                                  # 40 independent chroma blocks, recorded first
        c.append(cmd("IPRED", 1, 4, 4, 4 * (i % 8), 4 * (i // 8), flags=F_HEAD, u=u_ipred(4 * (i % 8), 4 * (i // 8), have=0)))
    c += wavefront_frame(rows=7, cols=8, chroma_free=0)  # then luma blocks: one dependent run, cut into units of 64 commands
    order, deps = tickets(c, est_weight=1)
    linked = set(u for u, d in enumerate(deps) if d) | set(p for d in deps for p in d)
    assert linked, "the fixture must contain dependent units"
    free = [u for u in range(len(order)) if u not in linked]
    assert len(free) >= 40
    # the head of the chains holds ticket 0, and no independent unit is queued before a unit that still
    # has more work behind it than in front of it
    assert order[0] in linked
    producers = set(p for d in deps for p in d)
    heads = [u for u in producers if not deps[u]]
    assert max(order.index(u) for u in heads) < min(order.index(u) for u in free)
