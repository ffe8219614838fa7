"""rtc_render_accumulate_static: the allocation-free variant of rtc_render_accumulate (wavefront scratch in a static
device array of abi.RTC_STATIC_SCRATCH_BYTES inside librtc.so; no counterpart in the reference, include/rtc.h).
Through the C ABI, against the existing entry point on the same inputs: the accumulators (and the resolved radiance /
RGB8) are BIT-IDENTICAL — same kernels, same order, only the scratch address differs — on calls in a row that share
the array (different scene sizes, seeds, structures, so every call runs over the previous call's leftovers), sizes
beyond the array are refused with nothing rendered, and device free memory does not move across the calls.
Collected last on purpose: a failure here hides no other test."""
import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc, scenes

pytestmark = pytest.mark.gpu


def _bits(t):
    return t.cpu().numpy().view(np.uint32)


def _cap(max_bounce):
    usable = abi.RTC_STATIC_SCRATCH_BYTES - 2 * (64 + 1) * 4 - 13 * 256
    return usable // (152 + 24 * max_bounce)


def test_static_scratch_variant_equals_render_accumulate():
    import torch
    dev = torch.device("cuda", 0)
    atrium = scenes.atrium(detail=0.35)
    cornell = scenes.cornell(subdiv=3)
    sc_a = rtc.Scene(rtc.Geometry(atrium.positions, atrium.tri_data, (64, 64, 64)), atrium.materials, device=0)
    sc_c = rtc.Scene(rtc.Geometry(cornell.positions, cornell.tri_data, (16, 16, 16)), cornell.materials, device=0)
    try:
        sc_a.build_bvh()
        cam_a = atrium.camera(height=90)          # 160 x 90
        cam_c = cornell.camera(height=64)
        # calls in a row on the shared array: grid then BVH on one scene, another scene with a deeper path, a band of a
        # frame with a sample offset (how a caller splits a frame that is too large for one call), the primary slice
        cases = [
            (sc_a, cam_a, dict(num_samples=6, max_bounce=4, seed=3)),
            (sc_a, cam_a, dict(num_samples=5, max_bounce=4, seed=9, accel=abi.RTC_ACCEL_BVH)),
            (sc_c, cam_c, dict(num_samples=7, max_bounce=9, seed=5)),
            (sc_a, cam_a, dict(num_samples=3, max_bounce=4, seed=3, sample_offset=6, total_samples=9, tile_rank=1, tile_world=3, tile_rows=8)),
            (sc_a, cam_a, dict(mode=abi.RTC_MODE_PRIMARY)),
            (sc_a, cam_a, dict(num_samples=6, max_bounce=4, seed=3, math=abi.RTC_MATH_FAST)),
        ]
        want = []
        for scene, cam, kw in cases:               # the existing entry point first (it allocates its own scratch)
            acc = torch.zeros(cam.w * cam.h, 4, device=dev)
            scene.render_accumulate(cam, rtc.make_params(**kw), acc.data_ptr())
            torch.cuda.synchronize()
            want.append(acc)
            assert float(acc[:, 3].max()) > 0.0

        def run_static():
            got = []
            for scene, cam, kw in cases:
                acc = torch.zeros(cam.w * cam.h, 4, device=dev)
                scene.render_accumulate_static(cam, rtc.make_params(**kw), acc.data_ptr())
                got.append(acc)
            torch.cuda.synchronize()
            return got

        for rep in range(2):                        # two rounds in a row over the same array
            got = run_static()
            for k, (g, w) in enumerate(zip(got, want)):
                assert np.array_equal(_bits(g), _bits(w)), "case %d, round %d" % (k, rep)
        # resolved outputs of the first case, through rtc_resolve: identical radiance and RGB8
        n = cam_a.w * cam_a.h
        outs = []
        for acc in (want[0], got[0]):
            rad = torch.zeros(n, 3, device=dev)
            rgb = torch.zeros(n, 3, dtype=torch.uint8, device=dev)
            rtc.resolve(acc.data_ptr(), n, 6, rgb.data_ptr(), rad.data_ptr(), device=0)
            torch.cuda.synchronize()
            outs.append((rad, rgb))
        assert np.array_equal(_bits(outs[0][0]), _bits(outs[1][0])) and torch.equal(outs[0][1], outs[1][1])

        # no device memory is taken by the calls (everything is warm by now: modules loaded, torch's pool holds the buffers)
        bufs = [torch.zeros_like(w) for w in want]
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        for (scene, cam, kw), acc in zip(cases, bufs):
            scene.render_accumulate_static(cam, rtc.make_params(**kw), acc.data_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(0)[0] == free0
        for b, w in zip(bufs, want):
            assert np.array_equal(_bits(b), _bits(w))

        # beyond the array: refused with a clear error, the accumulator untouched
        big = atrium.camera(height=270)             # 480 x 270 x 4 samples = 518 400 paths > 270 584
        assert big.w * big.h * 4 > _cap(4) >= cam_a.w * cam_a.h * 6
        acc = torch.zeros(big.w * big.h, 4, device=dev)
        with pytest.raises(rtc.RtcError) as e:
            sc_a.render_accumulate_static(big, rtc.make_params(num_samples=4, max_bounce=4), acc.data_ptr())
        assert e.value.code == abi.RTC_ERR_UNSUPPORTED and "static scratch" in str(e.value)
        with pytest.raises(rtc.RtcError) as e:      # 160 x 90 x 4 fits at max_bounce 4 but not at 64
            sc_a.render_accumulate_static(cam_a, rtc.make_params(num_samples=4, max_bounce=64), acc.data_ptr())
        assert e.value.code == abi.RTC_ERR_UNSUPPORTED
        with pytest.raises(rtc.RtcError) as e:      # the BVH build allocates: it must have been asked for beforehand
            sc_c.render_accumulate_static(cam_c, rtc.make_params(num_samples=2, accel=abi.RTC_ACCEL_BVH), acc.data_ptr())
        assert e.value.code == abi.RTC_ERR_INVALID_ARGUMENT
        with pytest.raises(rtc.RtcError):
            sc_a.render_accumulate_static(cam_a, rtc.make_params(num_samples=2), 0)
        torch.cuda.synchronize()
        assert float(acc.abs().max()) == 0.0
        # the same large frame in bands that fit: equal to the existing entry point on the whole frame
        ref = torch.zeros(big.w * big.h, 4, device=dev)
        sc_a.render_accumulate(big, rtc.make_params(num_samples=2, max_bounce=4, seed=11), ref.data_ptr())
        for r in range(4):
            sc_a.render_accumulate_static(big, rtc.make_params(num_samples=2, max_bounce=4, seed=11, tile_rank=r, tile_world=4), acc.data_ptr())
        torch.cuda.synchronize()
        assert np.array_equal(_bits(acc), _bits(ref))
    finally:
        sc_a.close()
        sc_c.close()
