"""Synthetic stress scenes (BASELINE config 5): an M-triangle random mesh + N random spheres, as a flattened
Scene whose tree comes from the bvh.h-equivalent builder (host.build_bvh).  Deterministic in (n_tri, n_sph, seed).

  triangles : centre ~ U([-50,50]^3), vertices = centre + U([-0.15,0.15]^3), UVs ~ U([0,1]^2), random winding (half are
              back-face culled by the reference's test, by design), 64 factor-only PBR materials
  spheres   : centre ~ U([-50,50]^3), r ~ U(0.05,0.3); 80 % PBR / 15 % metal (fuzz U(0,0.5)) / 5 % dielectric(1.5)
              (the ratios of main.cpp:100-118); + 16 emissive spheres r = 2, radiance 30
  camera    : eye (0,0,140) -> origin, vfov 45, aperture 0; 1920x1080, 256 spp, depth 16
"""
import numpy as np

from . import host
from .scene import Scene

F, I = np.float32, np.int32
PBR, METAL, DIEL, LIGHT = 0, 1, 2, 3


def make_stress_scene(n_tri=10_000_000, n_sph=100_000, seed=0xB200, width=1920, height=1080, spp=256, depth=16,
                      extent=50.0, threads=0, build=True):
    rng = np.random.default_rng(seed)
    s = Scene()
    s.name = "c5_stress_%dtri_%dsph" % (n_tri, n_sph)
    # materials: 64 PBR, 32 metal, 1 dielectric, 1 light
    n_pbr, n_metal = 64, 32
    kind = np.array([PBR] * n_pbr + [METAL] * n_metal + [DIEL, LIGHT], I)
    nm = len(kind)
    albedo = np.ones((nm, 4), F)
    albedo[:n_pbr + n_metal, :3] = rng.uniform(0.2, 0.9, (n_pbr + n_metal, 3)).astype(F)
    params = np.zeros((nm, 4), F)                      # metalness, roughness, fuzz, ir
    params[:n_pbr, 0] = (rng.random(n_pbr) < 0.5).astype(F)
    params[:n_pbr, 1] = rng.uniform(0.1, 0.9, n_pbr).astype(F)
    params[n_pbr:n_pbr + n_metal, 2] = rng.uniform(0.0, 0.5, n_metal).astype(F)
    params[n_pbr + n_metal, 3] = 1.5
    tex = np.full((nm, 4), -1, I)
    tex[nm - 1, 0] = 0                                 # the light's emit texture
    s.mat_kind, s.mat_tex, s.mat_albedo, s.mat_params = kind, tex, albedo, params
    s.tex_kind = np.array([0], I)                      # one solid colour: radiance 30
    s.tex_color = np.array([[30.0, 30.0, 30.0]], F)
    s.tex_evenodd = np.full((1, 2), -1, I)
    s.tex_whb = np.zeros((1, 3), I)
    s.tex_missing = np.zeros(1, I)
    # triangles
    c = rng.uniform(-extent, extent, (n_tri, 3)).astype(F)
    v = [(c + rng.uniform(-0.15, 0.15, (n_tri, 3)).astype(F)).astype(F) for _ in range(3)]
    flip = rng.random(n_tri) < 0.5
    v1 = np.where(flip[:, None], v[2], v[1]); v2 = np.where(flip[:, None], v[1], v[2])
    s.tri_v0, s.tri_v1, s.tri_v2 = v[0], np.ascontiguousarray(v1), np.ascontiguousarray(v2)
    s.tri_uv0, s.tri_uv1, s.tri_uv2 = (rng.random((n_tri, 2)).astype(F) for _ in range(3))
    s.tri_mat = rng.integers(0, n_pbr, n_tri).astype(I)
    del c, v, v1, v2
    # spheres (+ 16 lights)
    n_light = 16
    cs = rng.uniform(-extent, extent, (n_sph, 3)).astype(F)
    rs = rng.uniform(0.05, 0.3, n_sph).astype(F)
    u = rng.random(n_sph)
    ms = np.where(u < 0.8, rng.integers(0, n_pbr, n_sph),
                  np.where(u < 0.95, n_pbr + rng.integers(0, n_metal, n_sph), n_pbr + n_metal)).astype(I)
    cl = rng.uniform(-extent, extent, (n_light, 3)).astype(F)
    s.sph_c0 = np.concatenate([cs, cl]); s.sph_c1 = s.sph_c0.copy()
    s.sph_r = np.concatenate([rs, np.full(n_light, 2.0, F)])
    s.sph_t0 = np.zeros(n_sph + n_light, F); s.sph_t1 = np.ones(n_sph + n_light, F)
    s.sph_mat = np.concatenate([ms, np.full(n_light, nm - 1, I)])
    # camera + config
    s.cam_args = np.array([0, 0, 140, 0, 0, 0, 0, 1, 0, 45.0, width / height, 0.0, 10.0, 0.0, 1.0], F)
    a = s.cam_args
    s.camera = host.make_camera(a[0:3], a[3:6], a[6:9], *[float(x) for x in a[9:15]])
    s.render = np.array([width, height, spp, depth], I)
    if build:
        s.nodes, _ = host.build_bvh(s, 0.0, 1.0, rng_skip=0, threads=threads)
    return s
