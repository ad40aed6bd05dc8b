"""The parity cases shared by the golden generator, the CPU (emulator) tests and the GPU tests.
Each case = (name, scene builder, kwargs); every builder only talks to a yafrayInterface_t-shaped object."""
import numpy as np

from core_b200 import scenes

_RG_BLEND = dict(
    extra_materials=[
        ("rg", dict(type="rough_glass", IOR=1.5, alpha=0.3, filter_color=(0.8, 1.0, 0.9, 1.0), transmit_filter=0.7, fake_shadows=True)),
        ("rg2", dict(type="rough_glass", IOR=1.33, alpha=0.08, mirror_color=(1.0, 0.9, 0.8, 1.0))),
        ("bl_gs", dict(type="blend_mat", material1="glossy", material2="shiny", blend_value=0.35)),
        ("bl_gc", dict(type="blend_mat", material1="glass", material2="coated", blend_value=0.6)),
        ("bl_md", dict(type="blend_mat", material1="gdiff", material2="mirror", blend_value=0.75))],
    order=["rg", "bl_gs", "coated", "bl_gc", "gdiff", "rg2", "bl_md", "shiny"])

_MESHLIGHT = dict(
    emitters=[dict(name="emit1", id=701, centre=(-1.0, -2.5, 3.6), radius=0.5, subdiv=1, scale=(1.6, 1.0, 0.4),
                   material=dict(type="light_mat", color=(1.0, 0.8, 0.6, 1.0), power=6.0)),
              dict(name="emit2", id=702, centre=(3.0, 1.0, 3.0), radius=0.35, subdiv=0,
                   material=dict(type="light_mat", color=(0.6, 0.8, 1.0, 1.0), power=9.0, double_sided=True))],
    lights=[("ml1", dict(type="meshlight", object=701, color=(1.0, 0.8, 0.6, 1.0), power=6.0, samples=3)),
            ("ml2", dict(type="meshlight", object=702, color=(0.6, 0.8, 1.0, 1.0), power=9.0, samples=2, double_sided=True)),
            ("pt", dict(type="pointlight", color=(0.9, 0.9, 1.0, 1.0), power=6.0, **{"from": (-4.0, -3.0, 5.0)}))])

def _mapper(name, tex, texco="orco", scale=(1.0, 1.0, 1.0), offset=(0.0, 0.0, 0.0), **kw):
    return dict(name=name, type="texture_mapper", texture=tex, texco=texco, scale=scale, offset=offset, **kw)


# procedural textures (newperlin noise) through texture_mapper / value / mix / layer nodes on every node slot of the materials
_TEXTURED = dict(
    uv_orco=True,
    textures=[("tclouds", dict(type="clouds", depth=3, size=2.5, color1=(0.9, 0.3, 0.1, 1.0), color2=(0.1, 0.3, 0.9, 1.0))),
              ("tclouds_hard", dict(type="clouds", depth=2, size=4.0, hard=True, bias="negative")),
              ("tmarble", dict(type="marble", depth=2, size=1.5, turbulence=4.0, sharpness=2.5, shape="saw",
                               color1=(0.95, 0.95, 0.9, 1.0), color2=(0.2, 0.25, 0.3, 1.0))),
              ("twood", dict(type="wood", depth=2, size=3.0, turbulence=2.0, wood_type="rings", shape="tri",
                             color1=(0.55, 0.35, 0.15, 1.0), color2=(0.25, 0.12, 0.05, 1.0))),
              ("tcube", dict(type="rgb_cube")),
              ("tblend", dict(type="blend"))],
    extra_materials=[
        ("m_clouds", dict(type="shinydiffusemat", color=(1.0, 1.0, 1.0, 1.0), diffuse_reflect=0.9, diffuse_shader="map",
                          nodes=[_mapper("map", "tclouds", "orco", scale=(1.5, 1.5, 1.5))])),
        ("m_layer", dict(type="shinydiffusemat", color=(0.8, 0.8, 0.8, 1.0), diffuse_reflect=0.8, specular_reflect=0.1,
                         mirror_color=(1.0, 0.9, 0.8, 1.0), diffuse_shader="lay2", mirror_shader="mir", mirror_color_shader="mcol",
                         nodes=[_mapper("marb", "tmarble", "global", scale=(0.7, 0.7, 0.7), offset=(0.1, 0.2, 0.3)),
                                dict(name="lay1", type="layer", input="marb", mode=0, colfac=0.85, upper_color=(0.8, 0.2, 0.2, 1.0),
                                     upper_value=0.3, do_color=True, do_scalar=True, valfac=0.6, def_val=0.9),
                                _mapper("cl", "tclouds_hard", "orco", scale=(2.0, 2.0, 2.0)),
                                dict(name="lay2", type="layer", input="cl", upper_layer="lay1", mode=2, colfac=0.7, color_input=False,
                                     def_col=(0.2, 0.8, 0.3, 1.0), stencil=True, negative=True),
                                _mapper("mir", "twood", "orco", scale=(1.2, 1.2, 1.2)),
                                dict(name="mcol", type="value", color=(0.9, 0.8, 0.5, 1.0), alpha=1.0, scalar=0.4)])),
        ("m_glossy", dict(type="glossy", color=(0.9, 0.6, 0.3, 1.0), diffuse_color=(0.5, 0.2, 0.1, 1.0), diffuse_reflect=0.5,
                          glossy_reflect=0.7, exponent=60.0, as_diffuse=False, glossy_shader="gmix", diffuse_shader="dwood",
                          glossy_reflect_shader="grefl",
                          nodes=[_mapper("cube", "tcube", "uv"),
                                 dict(name="gmix", type="mix", mode=2, input1="cube", color2=(0.9, 0.8, 0.3, 1.0), value=0.6),
                                 _mapper("dwood", "twood", "global", scale=(0.5, 0.5, 0.5), proj_x=2, proj_y=3, proj_z=1),
                                 _mapper("grefl", "tblend", "uv")])),
        ("m_coated", dict(type="coated_glossy", color=(0.8, 0.2, 0.2, 1.0), diffuse_color=(0.6, 0.1, 0.1, 1.0), diffuse_reflect=0.5,
                          glossy_reflect=0.6, exponent=80.0, IOR=1.4, as_diffuse=False, diffuse_shader="ov",
                          nodes=[_mapper("a", "tmarble", "orco"), _mapper("b", "tclouds", "uv", scale=(3.0, 3.0, 1.0)),
                                 dict(name="ov", type="mix", mode=9, input1="a", input2="b", value=0.7)])),
        ("m_glass", dict(type="glass", IOR=1.5, filter_color=(0.8, 1.0, 0.9, 1.0), transmit_filter=0.6, fake_shadows=True,
                         mirror_color_shader="mc", nodes=[_mapper("mc", "tclouds", "orco", scale=(2.0, 2.0, 2.0))])),
        ("m_transp", dict(type="shinydiffusemat", color=(0.9, 0.7, 0.3, 1.0), diffuse_reflect=0.6, transparency=0.2, transmit_filter=0.7,
                          transparency_shader="tr", translucency_shader="tl", translucency=0.1,
                          nodes=[_mapper("tr", "tblend", "uv", scale=(2.0, 1.0, 1.0)),
                                 _mapper("c2", "tclouds_hard", "orco"),
                                 dict(name="tl", type="mix", mode=6, input1="c2", color2=(0.5, 0.5, 0.5, 1.0), factor="tr")]))],
    order=["m_clouds", "m_layer", "m_glossy", "m_glass", "m_coated", "m_transp", "mirror", "gdiff"])

# blend_mat with a mask shader node (blend.cc:49-66,406-427): the blend value is the scalar of a node graph at the shaded point -
# a texture_mapper over clouds, a marble texture through a layer node, and a constant value node
_BLEND_MASK = dict(
    uv_orco=True,
    textures=_TEXTURED["textures"],
    extra_materials=[
        ("bm_gs", dict(type="blend_mat", material1="glossy", material2="shiny", blend_value=0.35, mask="msk",
                       nodes=[_mapper("msk", "tclouds_hard", "orco", scale=(2.0, 2.0, 2.0))])),
        ("bm_gc", dict(type="blend_mat", material1="glass", material2="coated", blend_value=0.6, mask="lay",
                       nodes=[_mapper("marb", "tmarble", "global", scale=(0.7, 0.7, 0.7)),
                              dict(name="lay", type="layer", input="marb", mode=0, do_scalar=True, valfac=0.8, def_val=0.2)])),
        ("bm_md", dict(type="blend_mat", material1="gdiff", material2="mirror", blend_value=0.75, mask="val",
                       nodes=[dict(name="val", type="value", color=(0.5, 0.5, 0.5, 1.0), alpha=1.0, scalar=0.3)]))],
    order=["bm_gs", "glossy", "coated", "bm_gc", "gdiff", "shiny", "bm_md", "mirror"])

# glass / rough_glass with absorption (glass.cc:275-304, beer.cc): Beer's law over the segments inside the material - in the
# specular / glossy recursion (mcintegrator.cc:474-477,535-574,603-619) and along path segments (pathtracer.cc:290-293)
_ABSORB = dict(
    extra_materials=[
        ("ag", dict(type="glass", IOR=1.5, filter_color=(1.0, 1.0, 1.0, 1.0), transmit_filter=0.0, absorption=(0.9, 0.3, 0.2, 1.0),
                    absorption_dist=0.7, fake_shadows=True)),
        ("ag2", dict(type="glass", IOR=1.33, filter_color=(0.9, 1.0, 0.9, 1.0), transmit_filter=0.3, absorption=(0.2, 0.5, 0.95, 1.0))),
        ("arg", dict(type="rough_glass", IOR=1.5, alpha=0.2, absorption=(0.4, 0.9, 0.5, 1.0), absorption_dist=1.5))],
    order=["ag", "glossy", "arg", "ag2", "gdiff", "ag", "mirror", "arg"])

# the other noise bases of newNoise (blender = Blender's default basis, voronoi f1..crackle, cellnoise) under clouds / marble / wood
_NOISES = dict(
    uv_orco=True,
    textures=[("n_blender", dict(type="clouds", noise_type="blender", depth=3, size=3.0, color1=(0.9, 0.3, 0.1, 1.0), color2=(0.1, 0.3, 0.9, 1.0))),
              ("n_blender_hard", dict(type="marble", noise_type="blender", depth=2, size=2.0, turbulence=5.0, hard=True, sharpness=3.0, shape="tri")),
              ("n_vf1", dict(type="clouds", noise_type="voronoi_f1", depth=1, size=4.0)),
              ("n_vf2f1", dict(type="wood", noise_type="voronoi_f2f1", depth=1, size=3.0, turbulence=3.0)),
              ("n_crackle", dict(type="clouds", noise_type="voronoi_crackle", depth=0, size=5.0, color1=(1.0, 1.0, 0.8, 1.0), color2=(0.2, 0.1, 0.0, 1.0))),
              ("n_cell", dict(type="clouds", noise_type="cellnoise", depth=0, size=6.0)),
              ("n_vf3", dict(type="marble", noise_type="voronoi_f3", depth=1, size=2.0, turbulence=2.0)),
              ("n_vf4", dict(type="clouds", noise_type="voronoi_f4", depth=0, size=3.0, bias="positive")),
              ("n_std", dict(type="clouds", noise_type="stdperlin", depth=2, size=3.0, color1=(0.1, 0.5, 0.2, 1.0), color2=(1.0, 0.9, 0.7, 1.0)))],
    extra_materials=[(("mn%d" % i), dict(type="shinydiffusemat", color=(1.0, 1.0, 1.0, 1.0), diffuse_reflect=0.9, diffuse_shader="m",
                                         nodes=[_mapper("m", t, "orco" if i % 2 == 0 else "global", scale=(1.3, 1.3, 1.3))]))
                     for i, t in enumerate(["n_blender", "n_blender_hard", "n_vf1", "n_vf2f1", "n_crackle", "n_cell", "n_vf3", "n_std"])],
    order=["mn%d" % i for i in range(8)])

# voronoi (every distance metric and colour type), musgrave (five generators) and distorted-noise textures
_PROCS = dict(
    uv_orco=True,
    textures=[("v_int", dict(type="voronoi", size=3.0, weight1=1.0, weight2=0.5, intensity=1.2)),
              ("v_col1", dict(type="voronoi", size=4.0, color_type="col1", weight1=0.7, weight2=-0.3, weight3=0.2, distance_metric="manhattan")),
              ("v_col2", dict(type="voronoi", size=3.5, color_type="col2", weight1=1.0, weight4=0.4, distance_metric="chebychev")),
              ("v_col3", dict(type="voronoi", size=2.5, color_type="col3", weight1=1.0, weight2=1.0, distance_metric="minkovsky", mk_exponent=3.0)),
              ("v_mh", dict(type="voronoi", size=3.0, weight2=1.0, weight1=0.0, distance_metric="minkovsky_half")),
              ("v_m4", dict(type="voronoi", size=3.0, weight1=1.0, distance_metric="minkovsky_four", intensity=0.8)),
              ("m_fbm", dict(type="musgrave", musgrave_type="fBm", H=0.8, lacunarity=2.1, octaves=3.5, size=2.0, intensity=0.7,
                             color1=(0.1, 0.2, 0.5, 1.0), color2=(1.0, 0.9, 0.6, 1.0))),
              ("m_mf", dict(type="musgrave", musgrave_type="multifractal", noise_type="blender", H=0.9, lacunarity=2.0, octaves=3.0, size=1.5, intensity=0.5)),
              ("m_ht", dict(type="musgrave", musgrave_type="heteroterrain", H=1.0, lacunarity=2.0, octaves=2.5, offset=0.8, size=2.0, intensity=0.4)),
              ("m_hy", dict(type="musgrave", musgrave_type="hybridmf", noise_type="voronoi_f1", H=0.3, lacunarity=2.2, octaves=4.0, offset=0.7, gain=1.2,
                            size=1.5, intensity=0.6)),
              ("m_rd", dict(type="musgrave", musgrave_type="ridgedmf", H=1.0, lacunarity=2.0, octaves=4.0, offset=1.0, gain=2.0, size=1.5, intensity=0.5)),
              ("dn", dict(type="distorted_noise", noise_type1="blender", noise_type2="cellnoise", distort=1.5, size=3.0,
                          color1=(0.9, 0.9, 0.2, 1.0), color2=(0.2, 0.1, 0.6, 1.0))),
              ("dn2", dict(type="distorted_noise", distort=0.8, size=2.0))],
    extra_materials=[(("mp%d" % i), dict(type="shinydiffusemat", color=(1.0, 1.0, 1.0, 1.0), diffuse_reflect=0.9, diffuse_shader="m",
                                         nodes=[_mapper("m", t, "orco" if i % 2 == 0 else "global", scale=(1.2, 1.2, 1.2))]))
                     for i, t in enumerate(["v_int", "v_col1", "v_col2", "v_col3", "v_mh", "v_m4", "m_fbm", "m_mf", "m_ht", "m_hy", "m_rd", "dn", "dn2"])])

# bump mapping from procedural textures (textureMapper_t::evalDerivative, layer derivative chain, material_t::applyBump)
_BUMPED = dict(
    uv_orco=True,
    textures=_TEXTURED["textures"],
    extra_materials=[
        ("b_sd", dict(type="shinydiffusemat", color=(0.8, 0.7, 0.6, 1.0), diffuse_reflect=0.9, bump_shader="bmp", diffuse_shader="col",
                      nodes=[_mapper("bmp", "tclouds", "orco", scale=(3.0, 3.0, 3.0), bump_strength=4.0),
                             _mapper("col", "tmarble", "orco")])),
        ("b_glossy", dict(type="glossy", color=(0.9, 0.6, 0.3, 1.0), diffuse_color=(0.5, 0.2, 0.1, 1.0), diffuse_reflect=0.5,
                          glossy_reflect=0.7, exponent=60.0, as_diffuse=False, bump_shader="blay",
                          nodes=[_mapper("b1", "twood", "global", scale=(2.0, 2.0, 2.0), bump_strength=6.0),
                                 _mapper("b2", "tclouds_hard", "orco", scale=(2.0, 2.0, 2.0), bump_strength=3.0),
                                 dict(name="l1", type="layer", input="b1", mode=0),
                                 dict(name="blay", type="layer", input="b2", upper_layer="l1", mode=0, negative=True)])),
        ("b_coated", dict(type="coated_glossy", color=(0.8, 0.2, 0.2, 1.0), diffuse_color=(0.6, 0.1, 0.1, 1.0), diffuse_reflect=0.5,
                          glossy_reflect=0.6, exponent=80.0, IOR=1.4, as_diffuse=False, bump_shader="bm",
                          nodes=[_mapper("bm", "tmarble", "orco", scale=(1.5, 1.5, 1.5), bump_strength=5.0)])),
        ("b_glass", dict(type="glass", IOR=1.5, filter_color=(0.8, 1.0, 0.9, 1.0), transmit_filter=0.6, bump_shader="bg",
                         nodes=[_mapper("bg", "tclouds", "uv", scale=(4.0, 4.0, 1.0), bump_strength=2.0)]))],
    order=["b_sd", "b_glossy", "b_coated", "b_glass", "gdiff", "b_sd", "mirror", "b_glossy"])

# image textures (TGA files written by the scene builder): interpolation / clipping / repeat / crop modes, the cube / tube /
# sphere projections of texture_mapper, bump and normal maps from images
import os as _os
import tempfile as _tempfile
_TEXDIR = _os.path.join(_tempfile.gettempdir(), "yb200_tex")
_IMG_A, _IMG_B, _IMG_C = (_os.path.join(_TEXDIR, n) for n in ("a32.tga", "b24_top_rle.tga", "c8.tga"))
_IMAGED = dict(
    uv_orco=True,
    image_files=[(_IMG_A, dict(w=32, h=16, seed=5)), (_IMG_B, dict(w=20, h=24, seed=6, top=True, rle=True)),
                 (_IMG_C, dict(w=16, h=16, seed=7, gray=True))],
    textures=[("ia", dict(type="image", filename=_IMG_A, gamma=2.2)),
              ("ia_none", dict(type="image", filename=_IMG_A, interpolate="none", clipping="checker", even_tiles=True, odd_tiles=False,
                               checker_dist=0.2, xrepeat=1, yrepeat=1)),
              ("ib_cubic", dict(type="image", filename=_IMG_B, interpolate="bicubic", clipping="repeat", xrepeat=3, yrepeat=2, rot90=True)),
              ("ib_clip", dict(type="image", filename=_IMG_B, clipping="clip", cropmin_x=0.1, cropmax_x=0.8, cropmin_y=0.2, cropmax_y=0.9)),
              ("ic", dict(type="image", filename=_IMG_C, clipping="extend")),
              ("inorm", dict(type="image", filename=_IMG_A, normalmap=True))],
    extra_materials=[
        ("i_uv", dict(type="shinydiffusemat", color=(1.0, 1.0, 1.0, 1.0), diffuse_reflect=0.9, diffuse_shader="m",
                      nodes=[_mapper("m", "ia", "uv")])),
        ("i_checker", dict(type="shinydiffusemat", color=(0.8, 0.8, 0.8, 1.0), diffuse_reflect=0.9, diffuse_shader="lay",
                           nodes=[_mapper("m", "ia_none", "uv", scale=(4.0, 4.0, 1.0)),
                                  dict(name="lay", type="layer", input="m", mode=0, upper_color=(0.9, 0.9, 0.2, 1.0), use_alpha=True)])),
        ("i_cube", dict(type="glossy", color=(0.9, 0.6, 0.3, 1.0), diffuse_color=(0.5, 0.2, 0.1, 1.0), diffuse_reflect=0.6, glossy_reflect=0.5,
                        exponent=60.0, as_diffuse=False, diffuse_shader="m", bump_shader="bm",
                        nodes=[_mapper("m", "ib_cubic", "orco", mapping="cube"), _mapper("bm", "ic", "uv", scale=(2.0, 2.0, 1.0), bump_strength=2.0)])),
        ("i_tube", dict(type="shinydiffusemat", color=(0.8, 0.8, 0.8, 1.0), diffuse_reflect=0.9, diffuse_shader="m", transparency_shader="t",
                        transmit_filter=0.5, nodes=[_mapper("m", "ib_clip", "orco", mapping="tube"), _mapper("t", "ic", "orco", mapping="sphere")])),
        ("i_normal", dict(type="coated_glossy", color=(0.8, 0.2, 0.2, 1.0), diffuse_color=(0.6, 0.1, 0.1, 1.0), diffuse_reflect=0.5,
                          glossy_reflect=0.6, exponent=80.0, IOR=1.4, as_diffuse=False, bump_shader="nm",
                          nodes=[_mapper("nm", "inorm", "uv", bump_strength=0.5)]))],
    order=["i_uv", "i_checker", "i_cube", "i_tube", "i_normal", "i_uv", "mirror", "i_cube"])

IMAGE_CASES = {
    # BASELINE.json configs[0] (reduced): Cornell box, shinydiffuse + mirror component, area light, directlighting
    "c1_cornell_direct": (scenes.cornell_box, dict(width=64, height=64, spp=4, light_samples=4, mirror_box=True)),
    "c1_cornell_path": (scenes.cornell_box, dict(width=64, height=64, spp=4, light_samples=4, mirror_box=True,
                                                 integrator="pathtracing", integ_params=dict(bounces=4, path_samples=1))),
    "c1_cornell_gauss": (scenes.cornell_box, dict(width=48, height=48, spp=3, light_samples=2, mirror_box=False,
                                                  filter_type="gauss", pixelwidth=2.0)),
    # configs[1] (reduced): icosphere field, sun + point light, pathtracing 4 bounces
    "c2_field_path": (scenes.icosphere_field, dict(n_tris=20000, subdiv=2, width=96, height=54, spp=2, bounces=4)),
    "c2_field_flat_direct": (scenes.icosphere_field, dict(n_tris=6000, subdiv=1, width=64, height=36, spp=2, smooth=False,
                                                          integrator="directlighting")),
    # configs[2] (reduced): triangle soup, directlighting + ambient occlusion (IntersectS stress)
    "c3_soup_ao": (scenes.triangle_soup, dict(n_tris=3000, width=64, height=64, spp=1, ao_samples=8, ao_distance=0.3)),
    # configs[3] (reduced): glossy / coated glossy / glass, sun + point + area lights, pathtracing 8 bounces
    "c4_materials_path": (scenes.materials_scene, dict(width=80, height=45, spp=2, bounces=8)),
    "c4_materials_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting")),
    "c4_materials_tshadow": (scenes.materials_scene, dict(width=64, height=36, spp=1, integrator="directlighting",
                                                          integ_params=dict(transpShad=True, shadowDepth=4))),
    # the other cheap lights: directional (infinite and cylinder), spot (soft shadows: illumSample + MIS intersect), sphere
    "c4_lights_all_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="pathtracing", bounces=4, lights=[
        ("d", dict(type="directional", direction=(0.3, -0.5, 0.8), color=(1.0, 0.95, 0.9, 1.0), power=0.5)),
        ("s", dict(type="spotlight", to=(0.5, 0.5, 0.5), color=(1.0, 0.9, 0.8, 1.0), power=40.0, cone_angle=35.0, blend=0.2,
                   soft_shadows=True, samples=2, **{"from": (-3.0, -4.0, 6.0)})),
        ("sp", dict(type="spherelight", color=(0.8, 0.9, 1.0, 1.0), power=15.0, radius=0.6, samples=3, **{"from": (1.0, -3.0, 4.5)}))])),
    "c4_lights_dirac_direct": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="directlighting", lights=[
        ("d", dict(type="directional", direction=(0.1, -0.2, 0.9), color=(1.0, 1.0, 1.0, 1.0), power=1.5, infinite=False, radius=3.0,
                   **{"from": (0.5, -0.5, 6.0)})),
        ("s", dict(type="spotlight", to=(0.5, 0.5, 0.5), color=(1.0, 0.9, 0.8, 1.0), power=60.0, cone_angle=30.0, blend=0.3,
                   **{"from": (-3.0, -4.0, 6.0)}))])),
    # rough_glass (GGX refraction: two-direction sample of recursiveRaytrace) and blend_mat of two leaf materials
    "c4_roughglass_blend_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_RG_BLEND)),
    "c4_roughglass_blend_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=5, **_RG_BLEND)),
    # gradientBackground_t (gradientback.cc): sky / ground colours by ray direction, seen directly, through glass and in mirrors
    "c2_field_gradient_path": (scenes.icosphere_field, dict(n_tris=20000, width=96, height=54, spp=2, bounces=3,
                                                            background=dict(type="gradientback", horizon_color=(1.0, 0.9, 0.7, 1.0),
                                                                            zenith_color=(0.2, 0.4, 0.9, 1.0), power=1.3))),
    "c4_materials_gradient_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting",
                                                                  background=dict(type="gradientback", horizon_color=(0.9, 0.8, 0.6, 1.0),
                                                                                  zenith_color=(0.1, 0.3, 0.8, 1.0),
                                                                                  horizon_ground_color=(0.3, 0.25, 0.2, 1.0),
                                                                                  zenith_ground_color=(0.0, 0.0, 0.0, 1.0)))),
    "c4_absorb_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_ABSORB)),
    "c4_absorb_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=5, **_ABSORB)),
    "c4_blendmask_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_BLEND_MASK)),
    "c4_blendmask_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=4, **_BLEND_MASK)),
    "c4_blendmask_tshadow": (scenes.materials_scene, dict(width=64, height=36, spp=1, integrator="directlighting",
                                                          integ_params=dict(transpShad=True, shadowDepth=4), **_BLEND_MASK)),
    "c4_roughglass_blend_tshadow": (scenes.materials_scene, dict(width=64, height=36, spp=1, integrator="directlighting",
                                                                 integ_params=dict(transpShad=True, shadowDepth=4, bg_transp=True,
                                                                                   bg_transp_refract=True), **_RG_BLEND)),
    # meshlight bound to a visible emitter mesh (light_mat): area-cdf sampling + MIS intersect against the emitter triangles
    "c4_meshlight_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_MESHLIGHT)),
    # (path case on diffuse-only spheres: with glossy / specular recursion onto VISIBLE emitter geometry the reference's
    # state.includeLights leaks between sibling calls in traversal order - DESIGN.md section 7, first deviation)
    "c4_meshlight_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=4, order=["gdiff", "ground"] * 4, **_MESHLIGHT)),
    # shader nodes + procedural textures on UV / orco / global coordinates
    "c4_textured_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_TEXTURED)),
    "c4_textured_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=4, **_TEXTURED)),
    "c4_textured_tshadow": (scenes.materials_scene, dict(width=64, height=36, spp=1, integrator="directlighting",
                                                         integ_params=dict(transpShad=True, shadowDepth=4), **_TEXTURED)),
    "c4_noises_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_NOISES)),
    "c4_procs_a_direct": (scenes.materials_scene, dict(width=80, height=45, spp=1, integrator="directlighting",
                                                       order=["mp%d" % i for i in range(8)], **_PROCS)),
    "c4_procs_b_direct": (scenes.materials_scene, dict(width=80, height=45, spp=1, integrator="directlighting",
                                                       order=["mp%d" % i for i in (8, 9, 10, 11, 12, 0, 3, 6)], **_PROCS)),
    "c4_bump_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_BUMPED)),
    "c4_bump_path": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=3, **_BUMPED)),
    "c4_imagetex_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting", **_IMAGED)),
    "c4_imagetex_path": (scenes.materials_scene, dict(width=64, height=36, spp=1, bounces=3, **_IMAGED)),
    # depth of field with a polygonal lens (perspectiveCam_t::sampleTSD / biasDist)
    "c4_materials_dof_hexagon": (scenes.materials_scene, dict(width=64, height=36, spp=3, integrator="directlighting",
                                                              cam_params=dict(aperture=0.15, dof_distance=10.0, bokeh_type="hexagon",
                                                                              bokeh_rotation=15.0, bokeh_bias="center"))),
    # the other cameras: architect (perspective with a vertical image plane), orthographic, angular (fisheye; samples
    # outside the circle get camera weight 0)
    "c4_cam_architect": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="directlighting",
                                                      cam_params=dict(type="architect", focal=1.1))),
    "c4_cam_ortho": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="directlighting",
                                                  cam_params=dict(type="orthographic", scale=11.0))),
    "c4_cam_angular": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="pathtracing", bounces=3,
                                                    cam_params=dict(type="angular", angle=70.0, max_angle=60.0, circular=True,
                                                                    mirrored=True))),
    # adaptive anti-aliasing (AA_passes > 1: tiledIntegrator_t::render pass loop + imageFilm_t::nextPass)
    "c2_field_adaptive": (scenes.icosphere_field, dict(n_tris=6000, subdiv=1, width=64, height=36, spp=1, bounces=3, passes=3,
                                                       AA_inc_samples=3, AA_threshold=0.02)),
    "c4_materials_adaptive": (scenes.materials_scene, dict(width=80, height=45, spp=1, integrator="directlighting", passes=2,
                                                           AA_inc_samples=2, AA_threshold=0.1)),
    # scene_t::addInstance (rigid / uniform / non-uniform / sheared objToWorld) + angle-dependent smoothMesh
    "inst_smooth_direct": (scenes.instanced_scene, dict(width=96, height=54, spp=2)),
    "inst_smooth_path": (scenes.instanced_scene, dict(width=64, height=36, spp=2, integrator="pathtracing", bounces=3, smooth_angle=35.0)),
    # the other reconstruction filters (imagefilm.cc:119-165) and film options: premultiplied alpha, clamp_rgb, a cropped window
    "c1_cornell_mitchell": (scenes.cornell_box, dict(width=48, height=48, spp=3, light_samples=2, mirror_box=False,
                                                     filter_type="mitchell", pixelwidth=3.0)),
    "c1_cornell_lanczos": (scenes.cornell_box, dict(width=48, height=48, spp=3, light_samples=2, mirror_box=True,
                                                    filter_type="lanczos", pixelwidth=4.0)),
    "c1_cornell_gauss_wide": (scenes.cornell_box, dict(width=48, height=48, spp=2, light_samples=2, mirror_box=False,
                                                       filter_type="gauss", pixelwidth=6.0)),
    "c4_materials_premult_clamp": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="directlighting",
                                                                premult=True, clamp_rgb=True)),
    "c2_field_crop": (scenes.icosphere_field, dict(n_tris=6000, subdiv=1, width=40, height=24, spp=2, bounces=3,
                                                   xstart=17, ystart=9, cam_res=(96, 54))),
    # caustic photon map: mcIntegrator_t::createCausticMap / estimateCausticPhotons for both integrators - photons from
    # the sun, a point and an area light through glass / mirror / glossy spheres, gathered on the diffuse ground
    "c4_caustics_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting",
                                                        order=["glass", "mirror", "glass2", "coated", "gdiff", "glossy", "glass", "shiny"],
                                                        integ_params=dict(caustics=True, photons=30000, caustic_mix=40, caustic_depth=6,
                                                                          caustic_radius=0.35))),
    "c4_caustics_path_both": (scenes.materials_scene, dict(width=64, height=36, spp=2, bounces=4,
                                                           order=["glass", "mirror", "glass2", "coated", "gdiff", "glossy", "glass", "shiny"],
                                                           integ_params=dict(caustic_type="both", photons=20000, caustic_mix=30, caustic_depth=5,
                                                                             caustic_radius=0.4))),
    "c4_caustics_absorb_direct": (scenes.materials_scene, dict(width=80, height=45, spp=2, integrator="directlighting",
                                                               extra_materials=_ABSORB["extra_materials"],
                                                               order=["ag", "mirror", "ag2", "coated", "gdiff", "arg", "ag", "shiny"],
                                                               integ_params=dict(caustics=True, photons=30000, caustic_mix=40, caustic_depth=6,
                                                                                 caustic_radius=0.35))),
    "c4_caustics_lights_photon": (scenes.materials_scene, dict(width=64, height=36, spp=1, bounces=3, order=["glass", "mirror", "glass2", "gdiff"] * 2,
        integ_params=dict(caustic_type="photon", photons=24000, caustic_mix=25, caustic_depth=4, caustic_radius=0.5), lights=[
        ("d", dict(type="directional", direction=(0.3, -0.5, 0.8), color=(1.0, 0.95, 0.9, 1.0), power=0.6)),
        ("s", dict(type="spotlight", to=(0.5, 0.5, 0.5), color=(1.0, 0.9, 0.8, 1.0), power=60.0, cone_angle=35.0, blend=0.3,
                   **{"from": (-3.0, -4.0, 6.0)})),
        ("sp", dict(type="spherelight", color=(0.8, 0.9, 1.0, 1.0), power=25.0, radius=0.4, samples=2, **{"from": (1.0, -3.0, 4.5)}))])),
    # configs[4] (reduced): tessellated fBm terrain, sun + sky, pathtracing 4 bounces
    "c5_terrain_path": (scenes.terrain, dict(grid=60, width=64, height=36, spp=2, bounces=4)),
}

# Bars of the GPU image tests (mean relative error, PSNR dB, max alpha difference).  Default: the build rounds like the
# reference (-fmad=false), what is left is film summation order.  blend_mat samples wi = normalize(wi1 + wi2) (blend.cc:141):
# when the two sub-material directions nearly cancel, last-bit differences against the reference's -ffast-math build are
# amplified into a different path for single samples (measured: 5 pixels of 2304, rel 2.9e-6, PSNR 100.7 dB).
GPU_IMAGE_BAR = (1e-5, 100.0, 1e-5)
# Bump mapping differentiates the texture with a 2e-4 step (basicnodes.cc:27, 218-229): (f(x-h) - f(x+h)) / h turns the float
# rounding of f (about 6e-8) into about 3e-4 of the shading normal, so a -ffast-math build and a strict one cannot agree
# tighter than that; the north_star bar (1 % / 40 dB) still holds with a wide margin.
# Caustic photons through absorbing glass: exp(-sigma t) enters the photon's Russian roulette (material.cc:35-39); against the
# reference's -ffast-math build one photon of 30 000 ends elsewhere (one pixel off by 2 %, its gather neighbourhood by 1e-4:
# rel 5.2e-6, PSNR 95.6 dB).
GPU_IMAGE_BARS = {"c4_roughglass_blend_path": (1e-4, 90.0, 1e-4), "c4_bump_direct": (2e-3, 55.0, 1e-3), "c4_bump_path": (2e-3, 55.0, 1e-3),
                  "c4_caustics_absorb_direct": (2e-5, 90.0, 1e-5)}
EMU_IMAGE_BAR = (1e-4, 80.0, 1e-4)
EMU_IMAGE_BARS = {"c4_bump_direct": (2e-3, 55.0, 1e-3), "c4_bump_path": (2e-3, 55.0, 1e-3)}

# depth channel (z_channel): the frame AND the z values imageFilm_t::flush hands to colorOutput_t::putPixel
DEPTH_CASES = {
    # normalised, no far clip: range from tiledIntegrator_t::precalcDepths' per-pixel rays
    "z_materials_norm": (scenes.materials_scene, dict(width=64, height=36, spp=2, integrator="directlighting", z_channel=True)),
    # raw distances; misses read 99999997952
    "z_field_raw": (scenes.icosphere_field, dict(n_tris=6000, subdiv=1, width=64, height=36, spp=3, integrator="directlighting",
                                                 z_channel=True, normalize_z_channel=False, filter_type="gauss", pixelwidth=2.0)),
    # normalised against the camera's clip planes
    "z_materials_clip": (scenes.materials_scene, dict(width=48, height=27, spp=1, integrator="directlighting", z_channel=True,
                                                      cam_params=dict(nearClip=6.0, farClip=14.0))),
}

RAY_CASES = {
    "rays_cornell": (scenes.cornell_box, dict(width=32, height=32, spp=1)),
    "rays_field": (scenes.icosphere_field, dict(n_tris=20000, subdiv=2, width=64, height=36, spp=1)),
    "rays_soup": (scenes.triangle_soup, dict(n_tris=5000, width=32, height=32, spp=1)),
    "rays_terrain": (scenes.terrain, dict(grid=48, width=32, height=18, spp=1)),
}


def ray_batch(name, n=4096, seed=7):
    """Deterministic mixed ray set: directions from scene-dependent origins, finite and infinite tmax."""
    rng = np.random.Generator(np.random.PCG64(seed))
    o = rng.uniform(-1.0, 1.0, (n, 3))
    if name == "rays_field":
        o = o * np.array([8.0, 8.0, 1.0]) + np.array([0.0, 0.0, 2.0])
    if name == "rays_terrain":
        o = o * np.array([45.0, 45.0, 6.0]) + np.array([0.0, 0.0, 6.0])
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = o
    rays[:, 3] = np.where(rng.uniform(size=n) < 0.5, 0.0, 0.0005)
    rays[:, 4:7] = d
    rays[:, 7] = np.where(rng.uniform(size=n) < 0.5, -1.0, rng.uniform(0.2, 6.0, n))
    return rays


def compare_images(a, b):
    """mean relative error over RGB, PSNR (peak = max(1, ref max)), max alpha difference"""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    rel = np.abs(a[..., :3] - b[..., :3]).sum() / max(a[..., :3].sum(), 1e-12)
    mse = ((a[..., :3] - b[..., :3]) ** 2).mean()
    peak = max(1.0, a[..., :3].max())
    psnr = 10 * np.log10(peak * peak / max(mse, 1e-30))
    return rel, psnr, np.abs(a[..., 3] - b[..., 3]).max()


def compare_hits(ref, got, ties=None):
    """ref/got structured arrays (prim,t,b1,b2). Returns (id mismatch fraction excluding t-ties, max rel t err, max bary err)."""
    n = len(ref)
    same = ref["prim"] == got["prim"]
    both = (ref["prim"] >= 0) & (got["prim"] >= 0)
    tie = both & ~same & (np.abs(ref["t"] - got["t"]) <= 1e-5 * np.maximum(np.abs(ref["t"]), 1e-6))
    bad = ~same & ~tie
    ok = same & both
    terr = np.max(np.abs(ref["t"][ok] - got["t"][ok]) / np.maximum(np.abs(ref["t"][ok]), 1e-6)) if ok.any() else 0.0
    berr = max(np.max(np.abs(ref["b1"][ok] - got["b1"][ok])), np.max(np.abs(ref["b2"][ok] - got["b2"][ok]))) if ok.any() else 0.0
    return bad.sum() / max(n, 1), float(terr), float(berr), int(tie.sum())
