"""SPPM integrator: the device implementation against the oracle restatement of
integrator/photon_mapper.rs + light.rs.  The reference draws from thread_rng(), so parity is statistical: photon
statistics, the per-pixel SPPM state and the final image must agree within the oracle's own two-seed noise."""
import os

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from scenes_util import scene_builder, oracle_scene

W, H, SPP = 40, 40, 4
ITER, PHOTONS = 6, 20000


def _oracle(sb, seed, **kw):
    p = rt.copy_params(sb.params, width=W, height=H, spp_count=SPP, seed=seed)
    return oracle_scene("cornell_ref").sppm_render(sb.camera, p,  # the oracle's own scene.rs:16-112 graph
                                                    iterations=ITER, photons_per_iter=PHOTONS, **kw)


def test_oracle_sppm_known_answers(built):
    """Closed-form anchors of the restatement (no GPU): photon bookkeeping of generate_photon_map and SPPM::update."""
    sb = scene_builder("cornell_ref")
    acc, stats, (ng, nc) = _oracle(sb, 11)
    # every emitted photon leaves the ceiling light downwards into a closed box: its first hit is a surface, so at least
    # ~max(f) of them are stored; caustic photons need a specular hit first (the two spheres): a small, non-zero share
    assert ITER * PHOTONS * 0.5 < ng < ITER * PHOTONS * 8
    assert 0 < nc < 0.1 * ng
    g_n, c_n = stats[..., 4], stats[..., 9]
    seen = g_n > 0
    assert seen.mean() > 0.4  # the box fills about half of the 50-degree view from z = -800 (scene.rs:100-110)
    # photon_mapper.rs:51-52, 59: the count starts at 100 / 50 and only grows, by floor(0.7 * M)
    assert g_n[seen].min() >= 100 and c_n[seen].min() >= 50
    assert np.all(g_n[seen] == np.floor(g_n[seen])) and g_n[seen].max() > 100
    # radius2 only shrinks: after ITER rounds it is below the first k-nearest radius of a typical pixel, and positive
    assert np.all(stats[..., 3][seen] > 0)
    img = acc / SPP
    assert np.isfinite(img[seen]).mean() > 0.99


@pytest.mark.gpu
def test_sppm_matches_oracle(ctx):
    sb = scene_builder("cornell_ref")
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=W, height=H, spp_count=SPP, seed=5)
    acc, px, st = scene.render_sppm(sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS, want_pixel_stats=True)
    assert np.all(acc[..., 3] == SPP)
    a_acc, a_st, (a_ng, a_nc) = _oracle(sb, 101)
    b_acc, b_st, (b_ng, b_nc) = _oracle(sb, 202)
    print("photons gpu", st.global_photons, st.caustic_photons, "oracle", (a_ng, a_nc), (b_ng, b_nc), "ms", st.ms_total, st.ms_photons, st.ms_grid, st.ms_eye, st.ms_final)
    # 1. photon statistics (binomial counts over 120k emitted photons: sigma ~ 0.3 %; caustic ~ 2 %)
    ng, nc = 0.5 * (a_ng + b_ng), 0.5 * (a_nc + b_nc)
    assert abs(st.global_photons - ng) < 0.03 * ng  # seed-to-seed sigma is 0.5 % (8 seeds each side: +0.1 % +- 0.2 %)
    assert abs(st.caustic_photons - nc) < 0.10 * nc + 4 * abs(a_nc - b_nc)
    # 2. per-pixel SPPM state (photon_mapper.rs:33-64): photons, radius2 and flux, image means against the two oracle seeds
    for k, name, tol in ((4, "global photons", 0.03), (3, "global radius2", 0.08), (9, "caustic photons", 0.03), (8, "caustic radius2", 0.15)):
        g, a, b = float(np.mean(px[..., k])), float(np.mean(a_st[..., k])), float(np.mean(b_st[..., k]))
        ref = 0.5 * (a + b)
        print("  %-16s gpu %.5g oracle %.5g / %.5g" % (name, g, a, b))
        assert abs(g - ref) < tol * ref + 3 * abs(a - b), name
    for k0, name in ((0, "global flux"), (5, "caustic flux")):
        g = px[..., k0:k0 + 3].astype(np.float64).mean(axis=(0, 1))
        a, b = a_st[..., k0:k0 + 3].mean(axis=(0, 1)), b_st[..., k0:k0 + 3].mean(axis=(0, 1))
        ref = 0.5 * (a + b)
        print("  %-16s gpu %s oracle %s / %s" % (name, g, a, b))
        assert np.all(np.abs(g - ref) < 0.08 * ref + 3 * np.abs(a - b)), name
    # 3. the final image (photon_mapper.rs:327-365), NaN estimates (empty-map pixels) excluded on both sides
    gpu = acc[..., :3].astype(np.float64) / SPP
    a, b = np.nan_to_num(a_acc / SPP, nan=0.0, posinf=0.0), np.nan_to_num(b_acc / SPP, nan=0.0, posinf=0.0)
    ref = 0.5 * (a + b)
    cap = np.percentile(ref, 99.5) * 4 + 1e-3
    gm, rm = np.minimum(gpu, cap).mean(axis=(0, 1)), np.minimum(ref, cap).mean(axis=(0, 1))
    dab = np.minimum(a, cap) - np.minimum(b, cap)
    sigma_mean = np.sqrt((dab ** 2).mean(axis=(0, 1)) / 2.0 / (W * H))
    rel = np.abs(gm - rm) / np.maximum(rm, 1e-9)
    noise = sigma_mean * np.sqrt(1.5) / np.maximum(rm, 1e-9)
    print("  image mean gpu", gm, "oracle", rm, "rel", rel, "noise", noise)
    assert np.all(rel < np.maximum(0.05, 4 * noise))  # run-to-run sigma of the image mean is 1 % (8 seeds each side: -0.16 % +- 0.45 %)

    def tiles(img, t=8):
        return img[: H // t * t, : W // t * t].reshape(H // t, t, W // t, t, 3).mean(axis=(1, 3))
    tg, ta, tb = tiles(np.minimum(gpu, cap)), tiles(np.minimum(a, cap)), tiles(np.minimum(b, cap))
    env = np.sqrt(np.mean((ta - tb) ** 2)) / np.sqrt(2.0)
    err = np.sqrt(np.mean((tg - 0.5 * (ta + tb)) ** 2))
    print("  tile rms err %.5f, oracle two-seed sigma*sqrt(1.5) %.5f" % (err, env * np.sqrt(1.5)))
    assert err < 2.5 * env * np.sqrt(1.5) + 1e-3


@pytest.mark.gpu
def test_sppm_argument_errors(ctx):
    sb = scene_builder("json:scene_10")  # no marked light
    scene = rt.Scene(ctx, sb)
    p = rt.make_params(16, 16, 1)
    if sb.desc().n_lights == 0:
        with pytest.raises(rt.B200RTError) as e:
            scene.render_sppm(sb.camera, p, iterations=1, photons_per_iter=100)
        assert e.value.status == F.ERR_INVALID and "light" in str(e.value)
    sb2 = scene_builder("cornell_ref")
    scene2 = rt.Scene(ctx, sb2)
    for kw in (dict(iterations=0), dict(photons_per_iter=0), dict(k_global=129), dict(alpha=0.0)):
        args = dict(iterations=1, photons_per_iter=100)
        args.update(kw)
        with pytest.raises(rt.B200RTError):
            scene2.render_sppm(sb2.camera, rt.copy_params(sb2.params, width=16, height=16, spp_count=1), **args)


@pytest.mark.gpu
@pytest.mark.parametrize("n_ctx", [2, 3])
def test_sppm_multi_on_one_device_equals_render_sppm(ctx, n_ctx):
    """b200rt_render_sppm_multi with n contexts on ONE GPU (the driver's box): photons sharded by id, the photon maps gathered
    inside the grid build (plain local memory here, NVLink peer mappings on a multi-GPU box), pixel statistics sharded by pixel,
    final pass by sample index + k_reduce_peers.  Every draw is keyed by the global photon / pixel id, so the stored photon
    counts and the ray counts equal the one-device call's EXACTLY; statistics and image agree up to float summation order
    (the order of the photons inside a cell differs) and k-th-neighbour ties."""
    sb = scene_builder("cornell_ref")
    p = rt.copy_params(sb.params, width=W, height=H, spp_count=SPP + 1, seed=5)  # 5 samples over 2 or 3 shards: uneven split
    one = rt.Scene(ctx, sb)
    acc1, px1, st1 = one.render_sppm(sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS + 1, want_pixel_stats=True)
    scenes = [one] + [rt.Scene(rt.Context(0), sb) for _ in range(n_ctx - 1)]
    accm, pxm, stm = rt.render_sppm_multi(scenes, sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS + 1, want_pixel_stats=True)
    assert (stm.global_photons, stm.caustic_photons, stm.photon_rays, stm.eye_rays) == (st1.global_photons, st1.caustic_photons, st1.photon_rays, st1.eye_rays)
    assert np.all(accm[..., 3] == SPP + 1)
    assert np.array_equal(pxm[..., 4], px1[..., 4]) and np.array_equal(pxm[..., 9], px1[..., 9])  # photon counts per pixel: integers, order-independent
    for k in (3, 8):  # radius2: a product of exactly representable ratios applied in the same order
        np.testing.assert_allclose(pxm[..., k], px1[..., k], rtol=1e-5, atol=0)
    for k0 in (0, 5):  # flux: sums over the photons in range, in cell order
        np.testing.assert_allclose(pxm[..., k0:k0 + 3], px1[..., k0:k0 + 3], rtol=2e-3, atol=1e-3 * float(np.abs(px1[..., k0:k0 + 3]).max()))
    a, b = np.nan_to_num(accm[..., :3], nan=0.0, posinf=0.0), np.nan_to_num(acc1[..., :3], nan=0.0, posinf=0.0)
    assert abs(a.mean() - b.mean()) < 1e-3 * b.mean()
    assert np.mean(np.abs(a - b) <= 2e-3 * np.abs(b) + 1e-3 * b.mean()) > 0.999
    print("sppm multi x%d on one device: %.1f ms wall (one context: %.1f ms device)" % (n_ctx, stm.ms_total, st1.ms_total))


@pytest.mark.gpu
def test_sppm_multi_two_devices(ctx):
    import ctypes as C
    n = C.c_int(0)
    F.lib().b200rt_device_count(C.byref(n))
    if n.value < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    sb = scene_builder("cornell_ref")
    p = rt.copy_params(sb.params, width=W, height=H, spp_count=SPP, seed=5)
    s0, s1 = rt.Scene(ctx, sb), rt.Scene(rt.Context(1), sb)
    acc1, st1 = s0.render_sppm(sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS)
    accm, stm = rt.render_sppm_multi([s0, s1], sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS)
    assert (stm.global_photons, stm.caustic_photons, stm.photon_rays, stm.eye_rays) == (st1.global_photons, st1.caustic_photons, st1.photon_rays, st1.eye_rays)
    a, b = np.nan_to_num(accm[..., :3], nan=0.0, posinf=0.0), np.nan_to_num(acc1[..., :3], nan=0.0, posinf=0.0)
    assert abs(a.mean() - b.mean()) < 1e-3 * b.mean()


@pytest.mark.gpu
def test_sppm_multi_argument_errors(ctx):
    sb = scene_builder("cornell_ref")
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=16, height=16, spp_count=1)
    with pytest.raises(rt.B200RTError) as e:
        rt.render_sppm_multi([scene, scene], sb.camera, p, iterations=1, photons_per_iter=100)
    assert e.value.status == F.ERR_INVALID and "own context" in str(e.value)
    with pytest.raises(rt.B200RTError):
        rt.render_sppm_multi([scene], sb.camera, p, iterations=0, photons_per_iter=100)
    acc, st = rt.render_sppm_multi([scene], sb.camera, p, iterations=2, photons_per_iter=3000)  # one device: no peers at all
    assert st.global_photons > 1000 and np.all(acc[..., 3] == 1)


def _read_sppm_dump(path):
    raw = np.fromfile(path, dtype=np.uint8)
    n_g, n_c, n_pix, _ = np.frombuffer(raw[:16].tobytes(), dtype=np.uint32)
    f = np.frombuffer(raw[16:].tobytes(), dtype=np.float32).reshape(-1, 4)
    o = 0
    out = {}
    for name, n in (("g_pos", n_g), ("g_attr", 2 * n_g), ("c_pos", n_c), ("c_attr", 2 * n_c), ("vp", n_pix), ("vb", n_pix)):
        out[name] = f[o:o + int(n)]
        o += int(n)
    assert o == len(f)
    return out


def _d2(pos, p):
    """sppm_d2 of sppm.cuh: f32 subtractions, dx*dx rounded, then two fused multiply-adds (exact in f64, rounded once)."""
    d = (pos[None, :, :3] - p[:, None, :3]).astype(np.float32)
    t = (d[..., 0] * d[..., 0]).astype(np.float32)
    t = (d[..., 1].astype(np.float64) ** 2 + t.astype(np.float64)).astype(np.float32)
    return (d[..., 2].astype(np.float64) ** 2 + t.astype(np.float64)).astype(np.float32), d


def _flux(d, d2, sel, attr, bsdf):
    """sppm_photon_flux summed over the selected photons: bsdf * power * (1 - |n . d/|d||)"""
    power, nrm = attr[0::2, :3], attr[1::2, :3]
    dn = np.abs(np.einsum("pnk,nk->pn", d.astype(np.float64), nrm.astype(np.float64)) / np.sqrt(np.maximum(d2.astype(np.float64), 1e-300)))
    w = np.where(sel, 1.0 - dn, 0.0)
    return bsdf[:, :3].astype(np.float64) * (w @ power.astype(np.float64))


@pytest.mark.gpu
def test_sppm_lookups_match_brute_force(ctx, tmp_path):
    """The two look-ups of SPPM::update, pinned photon for photon (B200RT_SPPM_DUMP hands out every iteration's cell-ordered
    maps and visible points): the k-nearest start (PhotonMap::estimate_flux_by_count, photon_mapper.rs:77-96: radius^2 = the k-th
    smallest squared distance, flux over those k) of k_sppm_knn against a sort over ALL photons, and the radius gather of the next
    iteration (estimate_flux_within_radius, 99-114: squared distance < radius^2; N' = N + floor(alpha M), R'^2 = R^2 N' / (N + M),
    49-64) of k_sppm_gather against a count over ALL photons."""
    sb = scene_builder("cornell_ref")
    scene = rt.Scene(ctx, sb)
    w = h = 32
    p = rt.copy_params(sb.params, width=w, height=h, spp_count=1, seed=11)
    prefix = str(tmp_path / "sppm")
    os.environ["B200RT_SPPM_DUMP"] = prefix
    try:
        _, px1, _ = scene.render_sppm(sb.camera, p, iterations=1, photons_per_iter=20000, want_pixel_stats=True)
        _, px2, _ = scene.render_sppm(sb.camera, p, iterations=2, photons_per_iter=20000, want_pixel_stats=True)
    finally:
        del os.environ["B200RT_SPPM_DUMP"]
    px1, px2 = px1.reshape(-1, 10), px2.reshape(-1, 10)
    it0, it1 = _read_sppm_dump(prefix + ".it0.bin"), _read_sppm_dump(prefix + ".it1.bin")
    valid0 = it0["vp"][:, 3] != 0
    assert valid0.sum() > 0.3 * w * h  # the camera also sees past the box
    for col, kk, pos, attr in ((0, 100, it0["g_pos"], it0["g_attr"]), (5, 50, it0["c_pos"], it0["c_attr"])):
        assert len(pos) > kk
        d2, d = _d2(pos, it0["vp"][valid0])
        kth = np.partition(d2, kk - 1, axis=1)[:, kk - 1]
        got_r2 = px1[valid0, col + 3]
        assert np.array_equal(px1[valid0, col + 4], np.full(valid0.sum(), kk, np.float32))
        assert np.all(px1[~valid0, col + 4] == 0)
        np.testing.assert_allclose(got_r2, kth, rtol=3e-7, atol=0)  # the k-th nearest, to the last bit or the one next to it
        want = _flux(d, d2, d2 <= got_r2[:, None], attr, it0["vb"][valid0])
        np.testing.assert_allclose(px1[valid0, col:col + 3], want, rtol=2e-4, atol=1e-5 * np.abs(want).max())
    both = valid0 & (it1["vp"][:, 3] != 0)
    assert both.sum() > 0.3 * w * h
    for col, kk, pos, attr in ((0, 100, it1["g_pos"], it1["g_attr"]), (5, 50, it1["c_pos"], it1["c_attr"])):
        r2 = px1[both, col + 3]
        rr = (np.sqrt(r2).astype(np.float32) ** 2).astype(np.float32)  # r = sqrtf(radius2); squared distance < r * r
        d2, d = _d2(pos, it1["vp"][both])
        inside = d2 < rr[:, None]
        m = inside.sum(axis=1)
        n_new = kk + np.floor(0.7 * m.astype(np.float64)).astype(np.int64)
        off = px2[both, col + 4] != n_new
        assert off.sum() <= 2, (off.sum(), "pixels whose photon count differs (a distance on the radius, rounded the other way)")
        frac = (n_new.astype(np.float64) / (kk + m)).astype(np.float32)
        ok = ~off
        np.testing.assert_allclose(px2[both, col + 3][ok], (r2 * frac)[ok], rtol=3e-7, atol=0)
        want = (px1[both, col:col + 3].astype(np.float64) + _flux(d, d2, inside, attr, it1["vb"][both])) * frac[:, None]
        np.testing.assert_allclose(px2[both, col:col + 3][ok], want[ok], rtol=3e-4, atol=1e-5 * np.abs(want).max())


@pytest.mark.gpu
@pytest.mark.parametrize("photons,iters", [(40, 2), (300, 1), (5000, 9)])
def test_sppm_small_maps(ctx, photons, iters):
    """Edge shapes of the look-ups: fewer stored photons than k (the k-nearest start takes them all and the radius is the farthest
    one: kd_tree.nearests returns what there is), an empty or near-empty caustic map, a one-cell grid, more passes in flight than
    iterations, more iterations than photon buffers."""
    sb = scene_builder("cornell_ref")
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=24, height=20, spp_count=2, seed=9)
    acc, px, st = scene.render_sppm(sb.camera, p, iterations=iters, photons_per_iter=photons, want_pixel_stats=True)
    px = px.reshape(-1, 10)
    assert st.global_photons > 0 and st.photon_rays >= photons * iters
    seen = px[:, 4] > 0
    assert seen.sum() > 100
    assert np.all(px[seen, 4] >= 100) and np.all(px[seen, 9] >= 50)      # photons = k after the first look-up (photon_mapper.rs:51-52)
    assert np.all(np.isfinite(px[seen][:, [0, 1, 2, 3, 5, 6, 7, 8]])) and np.all(px[seen, 3] > 0)
    if photons * 3 < 100 * 2:  # fewer photons stored per iteration than k: the radius reaches the farthest photon of the first map
        assert np.all(px[seen, 3] > 1.0)
    assert np.all(acc[..., 3] == 2)


def _two_light_room(scale_right):
    """A closed white room with two equal XZ rectangle lights in the ceiling; the right one carries `scale_right` (light.rs:131)."""
    white, emit = rt.Lambertian(rt.ConstantTexture((0.73, 0.73, 0.73))), rt.DiffuseLight(rt.ConstantTexture((1.0, 1.0, 1.0)))
    left, right = rt.XZRectangle((100, 200), (200, 300), 554, emit), rt.XZRectangle((355, 200), (455, 300), 554, emit)
    walls = [rt.XZRectangle((0, 0), (555, 555), 0, white), rt.XZRectangle((0, 0), (555, 555), 555, white), rt.XYRectangle((0, 0), (555, 555), 555, white),
             rt.YZRectangle((0, 0), (555, 555), 0, white), rt.YZRectangle((0, 0), (555, 555), 555, white)]
    cam = rt.Camera(((278, 278, -800), (278, 278, 0)), (0, 1, 0), 40, 1, 0, 10)
    return rt.World(rt.HitableList(walls + [left, right]), cam, lights=[left, (right, scale_right)])


@pytest.mark.gpu
def test_sppm_lights_carry_their_own_scale(ctx):
    """Every upstream light has its own `scale` (SphereDiffuseLight::new / XZRectLight::new, light.rs:71,131): it weighs the choice
    between the lights (AllLights::emit, light.rs:214-219) and the power of the photons.  Two equal lights, the right one at 4x:
    the device against the oracle restatement (same tables), and against the equal-scale scene."""
    from oracle_ffi import OracleScene
    w = h = 40
    out = {}
    for name, sr in (("equal", 1.0), ("right4", 4.0)):
        world = _two_light_room(sr)
        sb = world.builder
        p = rt.copy_params(rt.make_params(w, h, 2), seed=3)
        scene = rt.Scene(ctx, sb)
        _, px, st = scene.render_sppm(world.cam.c, p, iterations=4, photons_per_iter=20000, light_scale=1e6, want_pixel_stats=True)
        flux = px[..., 0:3].astype(np.float64) / np.maximum(px[..., 3:4], 1e-30)  # flux / radius^2: proportional to irradiance
        out[name] = (flux, st)
        if name == "right4":
            o_flux = []
            for seed in (11, 12):
                _, o_px, _ = OracleScene(sb).sppm_render(world.cam.c, rt.copy_params(p, seed=seed), iterations=4, photons_per_iter=20000, light_scale=1e6)
                o_flux.append(o_px[..., 0:3] / np.maximum(o_px[..., 3:4], 1e-30))
            g, a, b = flux.mean(), o_flux[0].mean(), o_flux[1].mean()
            print("two lights, right at 4x: flux/r2 mean gpu %.5g oracle %.5g / %.5g" % (g, a, b))
            assert abs(g - 0.5 * (a + b)) < 0.06 * 0.5 * (a + b) + 3 * abs(a - b)
    eq, r4 = out["equal"][0], out["right4"][0]
    assert out["equal"][1].global_photons > 0
    # AllLights::emit picks a light with probability P_i / sum(P) and gives the photon that light's FULL power (light.rs:214-219:
    # no division by the pick probability), so the expected power per photon is sum(P_i^2) / sum(P): (16 + 1) / 5 = 3.4 against 1
    ratio_total = r4.mean() / eq.mean()
    assert 3.1 < ratio_total < 3.7, ratio_total
    # ... and one half of the image gains much more than the other (which one: Camera::new's u axis)
    halves = [r4[:, : w // 2].mean() / eq[:, : w // 2].mean(), r4[:, w // 2:].mean() / eq[:, w // 2:].mean()]
    assert max(halves) / min(halves) > 1.15, halves


def test_mark_light_scaled_rejects_bad_scales():
    b = rt.SceneBuilder()
    m = rt.DiffuseLight(rt.ConstantTexture((1, 1, 1))).flatten(b)
    prim = b._id(F.lib().b200rt_builder_rect(b.handle, 1, 0.0, 0.0, 1.0, 1.0, 5.0, m))
    for bad in (0.0, -1.0, float("nan"), float("inf")):
        assert F.lib().b200rt_builder_mark_light_scaled(b.handle, prim, bad) == F.ERR_INVALID
    assert F.lib().b200rt_builder_mark_light_scaled(b.handle, prim, 2.5) == 0
    d = b.desc()
    assert d.n_lights == 1 and abs(d.light_scales[0] - 2.5) < 1e-7


@pytest.mark.gpu
@pytest.mark.parametrize("gpus", [1, 2])
def test_cli_renders_sppm(ctx, tmp_path, gpus):
    """The reference's main() over the C ABI (main.rs:49-72: SPPMIntegrator::new + capture_image + save + the Total / SPPM / RT
    lines); with --gpus 2 through b200rt_render_sppm_multi."""
    import ctypes as C
    import subprocess
    from PIL import Image
    from rust_raytracer_b200.build import CLI
    n = C.c_int(0)
    F.lib().b200rt_device_count(C.byref(n))
    if n.value < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    out = str(tmp_path / "sppm.png")
    r = subprocess.run([CLI, "--scene", "cornell_ref", "--integrator", "sppm", "--width", "64", "--height", "64", "--spp", "4", "--iterations", "3", "--photons", "20000",
                        "--gpus", str(gpus), "--out", out], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Total:" in r.stdout and "SPPM:" in r.stdout and "RT:" in r.stdout and "%d GPU(s)" % gpus in r.stdout
    img = np.asarray(Image.open(out).convert("RGB"))
    assert img.shape == (64, 64, 3) and img.mean() > 5  # a lit Cornell box, not a black frame
