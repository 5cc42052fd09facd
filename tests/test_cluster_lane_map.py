"""The lane map of the row-resident inverse's grouped upper stage (csrc/acq_cluster.cuh, GeoW<8,8,25,20,2,512,0,1,4>), restated
as index arithmetic and checked on the CPU: every (column, pencil) of both passes is dealt to exactly one lane, every warp runs
three rounds, the busiest role of each group sits on a different scheduler, and every half-warp's sixteen 64-bit shared-memory
accesses — the two strided upper passes and the lower stage's stores into U — fall into sixteen different bank pairs with the
row pitch CW + 8."""
import itertools

RAH, RBH, CW, NW, UG, CPW = 25, 20, 32, 16, 4, 2
UP = CW + 8
KHW = 2 * UG


def lanes():
    for w, lane in itertools.product(range(NW), range(32)):
        grp = w // UG
        role = (w + grp) % UG
        yield w, lane, grp, role, role * 2 + (lane >> 4), lane & 7, (lane >> 3) & 1


def pass1_tiles(hw):
    return list(range(hw, RBH // 2, KHW))


def pass2_tiles(hw):
    return list(range((hw + KHW - 2) % KHW, (RAH + 1) // 2, KHW))


def test_every_pencil_is_dealt_exactly_once():
    seen1, seen2 = {}, {}
    for w, lane, grp, role, hw, c8, sub in lanes():
        col = grp * UG * CPW + c8
        for j in pass1_tiles(hw):
            seen1[(col, 2 * j + sub)] = seen1.get((col, 2 * j + sub), 0) + 1
        for j in pass2_tiles(hw):
            aa = 2 * j + sub
            if aa < RAH:
                seen2[(col, aa)] = seen2.get((col, aa), 0) + 1
    assert seen1 == {(c, ub): 1 for c in range(CW) for ub in range(RBH)}
    assert seen2 == {(c, aa): 1 for c in range(CW) for aa in range(RAH)}


def test_three_rounds_per_warp_and_roles_spread_over_the_schedulers():
    busiest = []
    for w in range(NW):
        grp = w // UG
        role = (w + grp) % UG
        r1 = max(len(pass1_tiles(role * 2 + h)) for h in (0, 1))
        r2 = max(len(pass2_tiles(role * 2 + h)) for h in (0, 1))
        assert r1 + r2 == 3, (w, r1, r2)               # one column per half-warp needed 2 + 2
        if r1 == 2:
            busiest.append(w % 4)                       # warp -> scheduler
    assert sorted(busiest) == [0, 1, 2, 3]              # one double radix-25 warp per group, each on another scheduler
    slots = sum(len(pass1_tiles(h)) + len(pass2_tiles(h)) for h in range(KHW)) * 16 * (NW // UG)
    used = CW * (RBH + RAH)
    assert used == 1440 and slots == 1472               # 98 % of the dealt lane slots carry a pencil (70 % before)


def _bank_pairs(addrs):
    return sorted(a % 16 for a in addrs)                # float2 elements: 16 bank pairs of 8 bytes


def test_upper_passes_and_lower_stores_are_bank_conflict_free():
    for w in range(NW):
        grp = w // UG
        role = (w + grp) % UG
        for half in (0, 1):
            hw = role * 2 + half
            half_lanes = [(l & 7, (l >> 3) & 1) for l in range(16)]
            for j in pass1_tiles(hw):
                for u in range(RAH):
                    addrs = [(RAH * (2 * j + sub) + u) * UP + grp * 8 + c8 for c8, sub in half_lanes]
                    assert _bank_pairs(addrs) == list(range(16))
            for j in pass2_tiles(hw):
                if 2 * j + 1 >= RAH:
                    continue                            # the half-full last tile: eight lanes, trivially distinct
                for u in range(RBH):
                    addrs = [(2 * j + sub + RAH * u) * UP + grp * 8 + c8 for c8, sub in half_lanes]
                    assert _bank_pairs(addrs) == list(range(16))
    # lower stage, pass B: lane = (block bl_l = lane / 8 of the warp's four, column aa = lane % 8); store b goes to U[gblk * UP + aa + 8 * (b % 4)]
    for gblk0 in range(0, 250, 4):
        for b in range(8):
            for half in (0, 1):
                addrs = [(gblk0 + 2 * half + bl) * UP + aa + 8 * (b % 4) for bl in (0, 1) for aa in range(8)]
                assert _bank_pairs(addrs) == list(range(16))
    # the old pitch (CW + 1) made those stores two-way conflicts
    addrs = [(bl) * (CW + 1) + aa for bl in (0, 1) for aa in range(8)]
    assert len(set(a % 16 for a in addrs)) < 16
