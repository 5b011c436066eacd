"""Pins the oracle (not gpu): the engine-side inputs the headless host restates are checked against the
reference's own known-answer arrays, and the unmodified ref_soft build reproduces the committed frames."""
import hashlib
import os

import numpy as np
import pytest

from conftest import EMU_PATH


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_gamma_tables_match_reference_goldens(pkg, scene, golden_gamma):
    """engine/client/gamma.c:245-268 Test_PrecomputedGammaTables, both parameter sets (:281, :357)."""
    m = scene.make_map("tiny", ncameras=16)
    r = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    for i in range(2):
        g, b, tg, lg = [float(x) for x in golden_gamma[f"params{i}"]]
        r.host.build_gamma(g, b, tg, lg)
        for which, name in enumerate(("texgammatable", "lightgammatable", "lineargammatable", "screengammatable")):
            got = r.host.gamma_table(which).astype(np.int64)
            want = golden_gamma[f"{name}{i}"].astype(np.int64)
            assert np.array_equal(got, want), (i, name)


@pytest.mark.parametrize("case", ["tiny_8x320x240", "small_16x320x240", "small_2x640x480"])
def test_oracle_reproduces_committed_frames(pkg, scene, golden_frames, case):
    meta, data = golden_frames
    c = meta["cases"][case]
    m = scene.make_map(c["map"], ncameras=max(c["views"], 16))
    assert hashlib.sha256(m.bsp).hexdigest() == c["bsp_sha256"], "map generator is not deterministic"
    r = scene.Renderer(pkg.REF_SOFT_PATH, c["width"], c["height"], m)
    col, z = r.render_views(m.cameras[: c["views"]])
    assert [sha(col[i]) for i in range(c["views"])] == c["colour_sha256"]
    assert [sha(z[i]) for i in range(c["views"])] == c["z_sha256"]
    for key, h in c["cache_blocks_sha256"].items():
        s, mip = map(int, key.split("."))
        assert sha(r.cache_block(s, mip)) == h, key
    if c["map"] == "tiny":
        assert np.array_equal(col, data["tiny_colour"]) and np.array_equal(z, data["tiny_z"])


def test_oracle_views_are_order_independent(pkg, scene):
    """The harness zeroes both buffers and forgets the cached view cluster per view, so a view's result
    does not depend on what was rendered before it (ref_soft never clears, r_edge.c:983)."""
    m = scene.make_map("tiny", ncameras=16)
    r = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    a, az = r.render_views(m.cameras[:6])
    b, bz = r.render_views(m.cameras[:6][::-1])
    assert np.array_equal(a, b[::-1]) and np.array_equal(az, bz[::-1])


@pytest.mark.parametrize("which,n,wh", [("tiny", 8, (320, 240)), ("small", 12, (320, 240)), ("small", 2, (640, 480))])
def test_stage_functions_match_oracle_on_cpu(pkg, scene, which, n, wh):
    """The scalar stage functions the kernels are built from (csrc/rc_core.h), run by the CPU debug harness:
    frames and z identical, every surface-cache block the oracle built is bit-exact."""
    if not os.path.exists(EMU_PATH):
        pytest.skip("debug harness not built")
    m = scene.make_map(which, ncameras=max(n, 16))
    so = scene.Renderer(pkg.REF_SOFT_PATH, *wh, m)
    em = scene.Renderer(EMU_PATH, *wh, m)
    rc, rz = so.render_views(m.cameras[:n])
    ec, ez = em.render_views(m.cameras[:n])
    assert (rc == ec).mean() >= 0.999 and (rz == ez).mean() >= 0.999
    ctr = em.counters()
    assert all(ctr[k] == 0 for k in ctr if k.startswith("overflow"))
    checked = 0
    for s in range(so.host.world_info()["numsurfaces"]):
        for mip in range(4):
            a = so.cache_block(s, mip)
            b = em.cache_block(s, mip)
            if a is not None and b is not None:
                assert np.array_equal(a, b), (s, mip)
                checked += 1
    assert checked > 0


@pytest.mark.parametrize("fmt", [(2, 0xF800, 0x07E0, 0x001F), (4, 0x00FF0000, 0x0000FF00, 0x000000FF), (4, 0x000000FF, 0x0000FF00, 0x00FF0000)])
def test_screen_map_restatement_matches_reference_blit(pkg, scene, fmt):
    """Display tail: what the unmodified R_EndFrame / R_BlitScreen writes to the software surface is the
    pixel_t frame looked up in the screen map; pins hostlib.screen_map (R_BuildScreenMap restated) for RGB565,
    XRGB8888 and XBGR8888 surfaces."""
    hostlib = __import__("importlib").import_module("xash3d-fwgs_b200.hostlib")
    m = scene.make_map("tiny", ncameras=16)
    r = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m, sw_format=fmt)
    col, _ = r.render_views(m.cameras[:4])
    disp = r.blit_views(m.cameras[:4])
    table = hostlib.screen_map(fmt)
    assert disp.dtype == (np.uint16 if fmt[0] == 2 else np.uint32)
    assert np.array_equal(disp.astype(np.uint32), table[col])
    assert len(np.unique(disp)) > 16


def test_scalar_triangle_rasteriser_matches_oracle_on_sprites(pkg, scene):
    """The scalar restatement of ref_soft's triangle rasteriser (csrc/rc_tri.h: final verts, 5-plane clip, edge tables,
    closed-form scanlines, span drawers) - the functions the entity kernels call per row and per pixel - driven
    sequentially by the CPU debug harness: sprites of every orientation type and four span drawers, bit-identical."""
    import importlib
    tg = importlib.import_module("test_gpu_parity")
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    for r in (emu, so):
        tg.add_test_sprites(r, m, m.cameras[32:64], (0, 1, 2, 3, 4))
    a, az = emu.render_views(m.cameras[:24])
    rc, rz = so.render_views(m.cameras[:24])
    assert np.array_equal(a, rc) and np.array_equal(az, rz)
    assert emu.counters()["tris"] > 0


@pytest.mark.parametrize("fastkeys", [False, True])
def test_scalar_studio_path_matches_oracle(pkg, scene, fastkeys, monkeypatch):
    """Studio models through the CPU debug harness: bones from the public library, the restated entity
    light, rc_studio_vertex / rc_studio_light per vertex and normal, then the triangle rasteriser.  fastkeys: the leading
    opaque run drawn as k_tris_fast / k_tri_resolve do (per-pixel maximum of z << 32 | submission number, one shade
    per pixel) - the order-free formulation must give R_PolysetFillSpans8's in-order result."""
    if fastkeys:
        monkeypatch.setenv("RC_EMU_FASTKEYS", "1")
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    bare, _ = so.render_views(m.cameras[:12])
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_studio_models(r, m.cameras, 24)
    a, az = emu.render_views(m.cameras[:12])
    rc, rz = so.render_views(m.cameras[:12])
    assert (rc != bare).mean() > 0.03, "the oracle drew no models"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)
    assert emu.counters()["tris"] > 1000


def test_light_gamma_change_reaches_the_stage_functions(pkg, scene):
    """same as the GPU test of the name, through the CPU debug harness: the host side resends the 2 KB table"""
    m = scene.make_map("tiny", ncameras=16)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    r0, _ = so.render_views(m.cameras[:4])
    emu.render_views(m.cameras[:4])
    for r in (emu, so):
        r.gamma_changed(gamma=2.2, brightness=0.0, texgamma=2.0, lightgamma=1.8)
    a1, az1 = emu.render_views(m.cameras[:4])
    r1, rz1 = so.render_views(m.cameras[:4])
    assert not np.array_equal(r1, r0)
    assert np.array_equal(a1, r1) and np.array_equal(az1, rz1)


def test_texture_dither_matches_oracle_on_cpu(pkg, scene):
    """sw_texfilt 1: rc_texfilt_index (the 2x2 dither kernel of D_DrawSpans16, r_scan.c:503-672) through the CPU harness"""
    m = scene.make_map("tiny", ncameras=16)
    emu = scene.Renderer(EMU_PATH, 320, 240, m, cvars={"sw_texfilt": 1})
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m, cvars={"sw_texfilt": 1})
    plain = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    a, az = emu.render_views(m.cameras[:4])
    rc, rz = so.render_views(m.cameras[:4])
    p, _ = plain.render_views(m.cameras[:4])
    assert not np.array_equal(p, rc), "sw_texfilt had no effect on the oracle"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)


@pytest.mark.parametrize("rotate", [False, True])
def test_scalar_brush_entity_path_matches_oracle(pkg, scene, rotate):
    """Opaque brush entities through the CPU debug harness: R_BmodelCheckBBox, R_FindTopnode, R_RotateBmodel,
    R_RecursiveClipBPoly / R_RenderBmodelFace and the leaf-keyed, 1/z-ordered surface stack of R_LeadingEdge, restated in
    csrc/rc_core.h - 20 boxes (doors / crates; rotated about all three axes in the second case) in 24 views."""
    m = scene.make_map("small", ncameras=64, brush_models=6)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    bare, _ = so.render_views(m.cameras[:24])
    for r in (emu, so):
        scene.add_brush_entities(r, m, m.cameras, 20, rotate=rotate)
    a, az = emu.render_views(m.cameras[:24])
    rc, rz = so.render_views(m.cameras[:24])
    assert (rc != bare).mean() > 0.1, "the oracle drew no brush entities"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)
    ctr = emu.counters()
    assert ctr["bmodel_surfs"] > 100 and ctr["overflow_bmodel"] == 0


def test_scalar_translucent_brush_entities_match_oracle(pkg, scene):
    """SURVEY 8a A12 through the CPU debug harness: R_DrawBrushModel in the sorted entity sequence - an edge frame per entity
    (rc_tbrush_pieces / rc_tbrush_line), D_BlendSpans16 / D_AddSpans16 / D_AlphaSpans16 pixels (rc_alpha_pixel) - next to
    opaque brush entities."""
    m = scene.make_map("small", ncameras=64, brush_models=6)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    opaque_only = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    scene.add_brush_entities(opaque_only, m, m.cameras, 8, rotate=True)
    base, _ = opaque_only.render_views(m.cameras[:28])
    for r in (emu, so):
        scene.add_mixed_brush_entities(r, m, m.cameras)
    a, az = emu.render_views(m.cameras[:28])
    rc, rz = so.render_views(m.cameras[:28])
    assert (rc != base).mean() > 0.05, "the oracle drew no translucent brush entities"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)
    assert emu.counters()["overflow_bmodel"] == 0


def test_scalar_chrome_meshes_match_oracle(pkg, scene):
    """STUDIO_NF_CHROME meshes (R_StudioSetupChrome, R_StudioDrawChromeMesh): texture coordinates that follow the viewer,
    rc_studio_chrome_st per (view, corner), through the CPU debug harness"""
    import importlib
    mdlgen = importlib.import_module("xash3d-fwgs_b200.tools.mdlgen")
    mdl = mdlgen.generate(nbones=12, ring=16, texflags=0x0002, skin=64)
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    plain = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_studio_models(r, m.cameras, 16, mdl=mdl)
    plain.set_time(1.0)
    scene.add_studio_models(plain, m.cameras, 16, mdl=mdlgen.generate(nbones=12, ring=16, texflags=0, skin=64))
    a, az = emu.render_views(m.cameras[:12])
    rc, rz = so.render_views(m.cameras[:12])
    p, _ = plain.render_views(m.cameras[:12])
    assert not np.array_equal(p, rc), "the chrome flag changed nothing in the oracle"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)


def test_light_point_slot_matches_oracle(pkg, scene):
    """the R_LightPoint slot (ref/common/ref_light.c:471) answers as ref_soft does at every camera position of the map"""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    got = [emu.light_point(c[:3]) for c in m.cameras[:64]]
    want = [so.light_point(c[:3]) for c in m.cameras[:64]]
    assert got == want and len(set(want)) > 8


def test_scalar_particles_match_oracle(pkg, scene):
    """CL_DrawParticles (ref/soft/r_part.c:47): particles as draw items whose view-facing quads are built per view
    (rc_particle_corners), additive drawer, palette colours, fading particles - through the CPU debug harness"""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    bare, _ = so.render_views(m.cameras[:12])
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_particles(r, m.cameras[:12], 240, ahead=40.0)
    a, az = emu.render_views(m.cameras[:12])
    rc, rz = so.render_views(m.cameras[:12])
    assert (rc != bare).mean() > 0.003, "the oracle drew no particles"
    assert np.array_equal(a, rc) and np.array_equal(az, rz)


def test_scalar_tracers_match_oracle(pkg, scene):
    """CL_DrawTracers (ref/soft/r_part.c:137): streaks whose quads are widened at right angles to their direction
    on the screen of every view (R_WorldToScreen through RI.worldviewProjectionMatrix, rc_view_wvp / rc_tracer_corners),
    every colour and width of the function's tables - through the CPU debug harness, with roll and other fields of view"""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = np.asarray(m.cameras[:12], dtype=np.float32).reshape(-1, 6).copy()
    cams[::3, 5] = 20.0  # roll
    bare, _ = so.render_views(cams)
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_tracers(r, cams, 96)
    for fov in (90.0, 70.0):
        a, az = emu.render_views(cams, fov_x=fov)
        rc, rz = so.render_views(cams, fov_x=fov)
        assert np.array_equal(a, rc) and np.array_equal(az, rz), ((a == rc).mean(), fov)
    assert (rc != bare).mean() > 0.002, "the oracle drew no tracers"


def test_tracer_list_moves_as_the_reference_moves_it(pkg, scene):
    """what CL_DrawTracers does to the client's list after drawing (r_part.c:236-253): positions advance by frametime,
    pt_grav tracers lose speed and fade, pt_slowgrav ones get their fall speed - one view, so both plugins run it once"""
    m = scene.make_map("tiny", ncameras=8)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    out = []
    for r in (emu, so):
        r.set_time(1.0)
        n = scene.add_tracers(r, m.cameras[:4], 40, moving=True)
        frame = r.render_views(m.cameras[:1])
        out.append((r.get_tracers(n), frame))
    assert np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[0][1][0], out[1][1][0])
    assert len(np.unique(out[1][0][:, 6])) > 4


def test_world_to_screen_slot_matches_oracle(pkg, scene):
    """the WorldToScreen slot (R_WorldToScreen, ref/soft/r_main.c:210) after a view: every camera position of the map seen from
    three views (fields of view 90 and 60, one rolled) answers as ref_soft does, bit for bit, behind-the-eye flag included"""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = np.asarray(m.cameras[:3], dtype=np.float32).reshape(-1, 6).copy()
    cams[2, 5] = 33.0
    for i, fov in enumerate((90.0, 60.0, 90.0)):
        got, want = [], []
        for r, acc in ((emu, got), (so, want)):
            r.render_views(cams[i:i + 1], fov_x=fov)
            acc.extend(r.world_to_screen(c[:3]) for c in m.cameras[:64])
        assert got == want
        assert 0 < sum(w[0] for w in want) < 64


def test_scalar_glow_sprites_match_oracle(pkg, scene):
    """kRenderGlow sprites (R_SpriteOccluded / R_SpriteGlowBlend, ref/soft/r_sprite.c:59-117): drawn or not, size and
    brightness per view - scissor on R_WorldToScreen's answer, the engine's sight line from every eye (here stopped by a wall
    the test places), distance falloff, kRenderFxNoDissipation - through the CPU debug harness"""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = np.asarray(m.cameras[:12], dtype=np.float32).reshape(-1, 6)
    bare, _ = so.render_views(cams)
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_glow_sprites(r, cams, 36)
    seen = []
    for axis, dist in ((-1, 0.0), (0, float(cams[:, 0].mean())), (1, float(cams[:, 1].mean()))):
        for r in (emu, so):
            r.set_sight_wall(axis, dist)
        a, az = emu.render_views(cams)
        rc, rz = so.render_views(cams)
        assert np.array_equal(a, rc) and np.array_equal(az, rz), ((a == rc).mean(), axis)
        seen.append(float((rc != bare).mean()))
    assert seen[0] > 0.002, "the oracle drew no glow"
    assert seen[1] < seen[0] or seen[2] < seen[0], "the wall hid nothing"


def test_scalar_decals_match_oracle(pkg, scene):
    """N2 decals (ref/soft/r_decals.c:719 R_DecalShoot, r_surf.c:914 R_DrawSurfaceDecals): decals shot at the walls the
    views look at - placement over the BSP, the clip of the face polygon against the decal, the blit into every mip of the
    face's cache blocks with the picture's 3-bit alpha - through the CPU debug harness.  The harness gives the oracle's faces
    their polygons (xh_model.c), without which ref_soft decides on an uninitialised count.  Pictures with an alpha plane only,
    as every decal of the game is: for an opaque one R_DrawSurfaceDecals reads `transparent` uninitialised (r_surf.c:932, :985).
    Scales >= 1 only: the blit ignores the decal's scale, and a decal placed with a smaller one can lie wholly above a face it was
    hung on, where the reference's unsigned row count wraps (r_surf.c:992-1000) and it runs off the block."""
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = m.cameras[:12]
    before, bz = so.render_views(cams)
    a0, _ = emu.render_views(cams)
    assert np.array_equal(a0, before)
    for r in (emu, so):
        scene.shoot_decals(r, cams, bz)
        scene.shoot_decals(r, cams[:8], bz, seed=8, scale=2.0)
    # ref_soft keeps using a block it cached before the shot unless the face is drawn in the very next frame (the shot sets
    # surf->dlightframe = framecount + 1, r_decals.c:538) or the block is rebuilt for another reason - which of the two depends on
    # its cache ring's history; ref/cuda rebuilds every block of the face at its next use.  The reference starts from an empty
    # cache here (R_GammaChanged -> D_FlushCaches) so that both show every decal.
    so.gamma_changed()
    after, az = so.render_views(cams)
    a, z = emu.render_views(cams)
    assert (after != before).mean() > 0.005, "the oracle drew no decals"
    assert np.array_equal(a, after) and np.array_equal(z, az), (a == after).mean()
    # other mip levels of the same faces: the views from further away / closer
    a, z = emu.render_views(m.cameras[12:40])
    rc, rz = so.render_views(m.cameras[12:40])
    assert np.array_equal(a, rc) and np.array_equal(z, rz), (a == rc).mean()
    # R_CreateDecalList: what a save game would hold
    la, lo = emu.decal_list(), so.decal_list()
    assert len(lo) >= 12 and np.array_equal(la, lo)
    # R_DecalRemoveAll: gone from the lists; blocks built from now on do not carry them
    for r in (emu, so):
        r.remove_decals(0)
        r.host.set_lightstyle(0, "z")
        r.host.run_lightstyles()
    a, z = emu.render_views(cams)
    rc, rz = so.render_views(cams)
    assert np.array_equal(a, rc) and np.array_equal(z, rz), (a == rc).mean()


def test_scalar_attachment_followers_match_oracle(pkg, scene):
    """R_StudioCalcAttachments (ref/soft/r_studio.c:1212) + the MOVETYPE_FOLLOW branch of R_DrawSpriteModel (r_sprite.c:170): sprites
    that ride on the attachment points of animated studio models are drawn where the model's bones put them this frame"""
    from importlib import import_module
    mdlgen = import_module("xash3d-fwgs_b200.tools.mdlgen")
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    mdl = mdlgen.generate(nbones=12, attachments=[(11, (6.0, 0.0, 4.0)), (5, (0.0, 9.0, 0.0)), (2, (-7.0, 3.0, 1.0))])
    res = []
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_studio_models(r, m.cameras, 6, nbones=12, mdl=mdl)
        spr = r.create_sprite("spr_follow", scene.sprite_image(17, 32, 32), sprtype=2, texformat=1)
        for k in range(6):
            # entity k + 1 is the k-th model; body = attachment number + 1 (0: the parent's origin)
            r.add_entity(spr, (0, 0, 0), rendermode=5, renderamt=255, rendercolor=(255, 255, 255), movetype=12, aiment=k + 1, body=k % 4, scale=0.5)
        res.append(r.render_views(m.cameras[:6]))
    bare = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    bare.set_time(1.0)
    scene.add_studio_models(bare, m.cameras, 6, nbones=12, mdl=mdl)
    nb, _ = bare.render_views(m.cameras[:6])
    assert (res[1][0] != nb).mean() > 0.001, "the oracle drew no follower"
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]), (res[0][0] == res[1][0]).mean()


def test_scalar_follower_models_match_oracle(pkg, scene):
    """models that ride on other models (MOVETYPE_FOLLOW, R_DrawStudioModel's branch r_studio.c:2869 + R_StudioMergeBones :780): the
    follower's bones that carry a parent bone's name take the parent's matrices, the others hang off them with its own animation"""
    from importlib import import_module
    mdlgen = import_module("xash3d-fwgs_b200.tools.mdlgen")
    m = scene.make_map("small", ncameras=64)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    parent = mdlgen.generate(nbones=12)
    rider = mdlgen.generate(nbones=6, ring=16, radius=14.0, seed=77, name="rider.mdl", bone_names=["bone08", "bone09", "w0", "w1", "w2", "w3"])
    res = []
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_studio_models(r, m.cameras, 6, nbones=12, mdl=parent)
        mi = r.load_studio("models/rider.mdl", rider)
        for k in range(6):
            r.add_entity(mi, (0, 0, 0), angles=(0.0, 11.0 * k, 0.0), frame=(5 * k) % 30, movetype=12, aiment=k + 1, animtime=0.4)
        res.append(r.render_views(m.cameras[:6]))
    bare = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    bare.set_time(1.0)
    scene.add_studio_models(bare, m.cameras, 6, nbones=12, mdl=parent)
    nb, _ = bare.render_views(m.cameras[:6])
    assert (res[1][0] != nb).mean() > 0.002, "the oracle drew no follower"
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]), (res[0][0] == res[1][0]).mean()


def test_scalar_decals_on_brush_entities_match_oracle(pkg, scene):
    """R_DecalShoot on brush entities (r_decals.c:736-779): the point goes into the entity's own space (rotated ones through the
    inverse of its matrix), the decal hangs on a face of the submodel and rides with it"""
    m = scene.make_map("small", ncameras=64, brush_models=4)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = np.asarray(m.cameras[:10], dtype=np.float32).reshape(-1, 6)
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_brush_entities(r, m, cams, 10, rotate=True)
    bare = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    nb, _ = bare.render_views(cams)
    before, bz = so.render_views(cams)
    # a pixel of every view that shows its entity: the middle of the pixels the entities changed
    spots = []
    for k in range(len(cams)):
        ys, xs = np.nonzero(before[k] != nb[k])
        spots.append((int(np.median(xs)), int(np.median(ys))) if len(xs) else (160, 120))
    rng = np.random.default_rng(3)
    img = np.zeros((32, 32, 4), np.uint8)
    img[..., 0] = 255; img[..., 1] = rng.integers(0, 255, (32, 32)); img[..., 3] = rng.integers(0, 255, (32, 32)); img[8:24, 8:24, 3] = 255
    for r in (emu, so):
        tex = r.load_pic("{entdecal", img, 1 << 16)
        for k, c in enumerate(cams):
            pt = scene.pixel_point(c, bz[k], spots[k][0], spots[k][1])
            r.decal_shoot_entity(tex, k + 1, pt, 0, 1.0)
    so.gamma_changed()
    rc, rz = so.render_views(cams)
    a, z = emu.render_views(cams)
    assert (rc != before).mean() > 0.001, "no decal landed on a brush entity"
    assert np.array_equal(a, rc) and np.array_equal(z, rz), (a == rc).mean()
    assert np.array_equal(emu.decal_list(), so.decal_list())


def test_scalar_coincident_edges_match_oracle(pkg, scene):
    """equal-u ties. Boxes whose faces lie in the floor's plane, touch each other and share edge lines
    (coincident edges of different surfaces, 1/z ties between pieces) seen from 24 cameras, some very close (edges clamped to
    the screen border): every pixel as ref_soft orders them"""
    m = scene.make_map("small", ncameras=64, brush_models=4)
    emu = scene.Renderer(EMU_PATH, 320, 240, m)
    so = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    cams = np.asarray(m.cameras[:24], dtype=np.float32).reshape(-1, 6).copy()
    for r in (emu, so):
        r.set_time(1.0)
        scene.add_abutting_boxes(r, m, cams[:12])
    near = cams.copy()  # views 12..23: the first twelve again from 60 units closer
    near[12:] = np.concatenate([cams[:12, :3] + 60.0 * np.stack([np.cos(np.radians(cams[:12, 4])), np.sin(np.radians(cams[:12, 4])), np.zeros(12)], -1), cams[:12, 3:]], -1)
    a, az = emu.render_views(near)
    rc, rz = so.render_views(near)
    bare = scene.Renderer(pkg.REF_SOFT_PATH, 320, 240, m)
    nb, _ = bare.render_views(near)
    assert (rc != nb).mean() > 0.05, "the boxes are not in view"
    assert np.array_equal(a, rc) and np.array_equal(az, rz), ((a == rc).mean(), (az == rz).mean())
