"""The kx sub-block pipeline of the y-slab pressure solve (multigpu.sub_block_plan, MHH_SLAB_A2A=ce_sub) as pure data
movement, on the CPU: every rank's x lines are scattered into segments as mhh_slab_fft_x_fwd_sub does (identity instead of the
transform), the transfers of the plan are carried out, the y side gathers as mhh_slab_fft_y_fwd_sub reads — and must see the
global field transposed as Transpose::exec_xy defines it (src/transpose.cxx:163-190: rank r receives the kx block r of every
rank's rows); then the way back (mhh_slab_fft_y_bwd_sub -> transfers -> mhh_slab_fft_x_bwd_sub) must restore every slab.
Layouts as declared in include/mhh_b200.h."""
import numpy as np
import pytest

from microhh_b200.multigpu import sub_block_count, sub_block_plan


@pytest.mark.parametrize("world,itot,jtot,kmax,wanted", [(1, 32, 8, 3, 4), (2, 32, 8, 3, 4), (4, 64, 16, 2, 4), (8, 256, 16, 2, 4), (2, 32, 8, 3, 2)])
def test_sub_block_pipeline_moves_every_value_to_its_place(world, itot, jtot, kmax, wanted):
    iblock, jmax = itot // world, jtot // world
    B = sub_block_count(iblock, world, wanted)
    assert iblock % B == 0 and world * B <= 32
    ib = iblock // B
    Rb = kmax * jmax * ib
    rng = np.random.default_rng(world * 100 + B)
    G = rng.standard_normal((kmax, jtot, itot))
    X = [G[:, r * jmax:(r + 1) * jmax, :].copy() for r in range(world)]                  # slab layout [k][j < jmax][i < itot]
    bufs = [dict(S=np.full(world * B * Rb, np.nan), R1=np.full(world * B * Rb, np.nan), R2=np.full(world * B * Rb, np.nan))
            for _ in range(world)]
    plans = [sub_block_plan(world, r, B) for r in range(world)]
    seg = lambda a, slot: a[slot * Rb:(slot + 1) * Rb].reshape(kmax, jmax, ib)

    # forward x pass: segment v of every line -> plan.x_fwd[v]
    for me in range(world):
        for v, (where, slot) in enumerate(plans[me]["x_fwd"]):
            seg(bufs[me][where], slot)[...] = X[me][:, :, v * ib:(v + 1) * ib]
    # forward transfers, sub-block by sub-block
    for b in range(B):
        for me in range(world):
            for r, dst, src in plans[me]["copies"][b]:
                seg(bufs[r]["R1"], dst)[...] = seg(bufs[me]["S"], src)
        # y side of sub-block b: element j = (r, j') of line (ii', k) at segment r * B + b
        for me in range(world):
            Yb = np.concatenate([seg(bufs[me]["R1"], r * B + b) for r in range(world)], axis=1)      # [k][j < jtot][ii < ib]
            i0 = me * iblock + b * ib
            assert np.array_equal(Yb, G[:, :, i0:i0 + ib]), (world, me, b)
            # the way back: rows of rank r -> plan.y_bwd[b][r]; the staging slots are free again (their forward copies are done)
            for r, (where, slot) in enumerate(plans[me]["y_bwd"][b]):
                seg(bufs[me][where], slot)[...] = Yb[:, r * jmax:(r + 1) * jmax, :]
        for me in range(world):
            for r, dst, src in plans[me]["copies"][b]:
                seg(bufs[r]["R2"], dst)[...] = seg(bufs[me]["S"], src)
    # inverse x pass: element m of a line from segment m // ib of R2
    for me in range(world):
        back = np.concatenate([seg(bufs[me]["R2"], v) for v in range(world * B)], axis=2)
        assert np.array_equal(back, X[me]), (world, me)


def test_sub_block_count_respects_the_segment_table_and_the_block():
    assert sub_block_count(128, 8) == 4          # 1024^3 on 8 ranks: 32 columns per sub-block, 32 table entries
    assert sub_block_count(256, 2) == 4
    assert sub_block_count(16, 2) == 2           # at least 8 columns per sub-block
    assert sub_block_count(8, 4) == 1
    assert sub_block_count(96, 2, 4) == 4 and sub_block_count(100, 2, 8) == 4
