"""Synthetic benchmark scenes (BASELINE.json configs 3-5; recipes in SURVEY.md section 8d).

All generators are deterministic functions of a SplitMix64 seed and return scene-description dicts (scene.py).
They only use materials, lights, cameras and backgrounds that exist upstream (src/scene.rs), so the oracle can
render the same description.
"""
import numpy as np

from . import scene as S

_MASK = (1 << 64) - 1

# BASELINE configs 1 and 2: the Cornell-box configuration of the upstream test_scene.txt, values restated (the file is not
# copied; tests/test_oracle_outbmp.py checks that it parses to the same scene as the original where that is present)
TEST_SCENE_TEXT = """
{
    objects: [
        { bounds: Plane { point: (0, 0, -3) normal: (0, 0, 1) }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 1.0, 1.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Plane { point: (0, 0, 0) normal: (0, 1.0, 0) }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 1.0, 1.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Plane { point: (0, 6, 0) normal: (0, -1.0, 0) }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 1.0, 1.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Plane { point: (-3, 0, 0) normal: (1, 0, 0) }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 0.0, 0.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Plane { point: (3, 0, 0) normal: (-1, 0, 0) }
          material: IndirectPhongMaterial { diffuse: rgb(0.0, 1.0, 0.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Sphere { center: (0, 1.5, 0) radius: 1.5 }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 1.0, 1.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(0.0, 0.0, 0.0) samples: 1 } }
        { bounds: Sphere { center: (0, 10.65, 0) radius: 5 }
          material: IndirectPhongMaterial { diffuse: rgb(1.0, 1.0, 1.0) specular: rgb(0.0, 0.0, 0.0) exponent: 1.0 ambient: rgb(5.0, 5.0, 5.0) samples: 1 } }
    ]
    lights: [ ]
    camera: SimplePerspectiveCamera new((0, 3, 17), (0, 0, -1), (0, 1, 0), 3.6)
    background: SolidColorBackground { color: rgb(0.051, 0.051, 0.051) }
    options: { width: 800 height: 800 antialias: 1024 }
}
"""


class SplitMix64:
    def __init__(self, seed):
        self.s = seed & _MASK

    def next_u64(self, n=1):
        """n successive outputs as a uint64 array (vectorised: state advances by the golden gamma per draw)."""
        idx = np.arange(1, n + 1, dtype=np.uint64)
        with np.errstate(over="ignore"):
            z = np.uint64(self.s) + idx * np.uint64(0x9E3779B97F4A7C15)
            self.s = int(z[-1])
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return z ^ (z >> np.uint64(31))

    def uniform(self, n, lo=0.0, hi=1.0):
        u = (self.next_u64(n) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
        return lo + (hi - lo) * u


def random_spheres(n, seed, extent, width, height, antialias=1, camera=None, transparent=True, reflective_only=False, background=None):
    """Ground plane + n random spheres.

    radius ~ U[0.2, 0.8]; centre x,z ~ U[-extent, extent], y = radius + U[0, 6];
    materials: 70 % Phong (diffuse U[0.1,0.9]^3, specular 0.3 grey, exponent 50, ambient 0.02*diffuse),
               20 % Fresnel (same + ior 1.5), 10 % Transparent (specular 0.9 grey, exponent 200, ior 1.5);
    reflective_only: 75 % Phong / 25 % Fresnel (config 5).
    Object 0 is the ground plane; spheres follow in generation order.
    """
    rng = SplitMix64(seed)
    radius = rng.uniform(n, 0.2, 0.8)
    cx = rng.uniform(n, -extent, extent)
    cz = rng.uniform(n, -extent, extent)
    cy = radius + rng.uniform(n, 0.0, 6.0)
    kind_u = rng.uniform(n)
    dr, dg, db = rng.uniform(n, 0.1, 0.9), rng.uniform(n, 0.1, 0.9), rng.uniform(n, 0.1, 0.9)
    b = S.SceneBuilder()
    ground = b.add_material(S.PhongMaterial((0.6, 0.6, 0.6), (0.2, 0.2, 0.2), 30.0, (0.012, 0.012, 0.012)))
    b.add_object(S.Plane((0.0, 0.0, 0.0), (0.0, 1.0, 0.0)), ground)
    mats = np.zeros(n, dtype=np.uint32)
    for i in range(n):
        d = (dr[i], dg[i], db[i])
        amb = (0.02 * dr[i], 0.02 * dg[i], 0.02 * db[i])
        u = kind_u[i]
        if reflective_only:
            m = S.PhongMaterial(d, (0.3, 0.3, 0.3), 50.0, amb) if u < 0.75 else S.FresnelMaterial(d, (0.3, 0.3, 0.3), 50.0, amb, 1.5)
        elif u < 0.7:
            m = S.PhongMaterial(d, (0.3, 0.3, 0.3), 50.0, amb)
        elif u < 0.9 or not transparent:
            m = S.FresnelMaterial(d, (0.3, 0.3, 0.3), 50.0, amb, 1.5)
        else:
            m = S.TransparentMaterial((0.9, 0.9, 0.9), 200.0, 1.5)
        mats[i] = b.add_material(m)
    b.add_spheres(np.stack([cx, cy, cz, radius], axis=1), mats)
    b.add_light(S.PointLight((0.5 * extent, 60.0, 0.6667 * extent), (1.0, 1.0, 1.0)))
    b.add_light(S.PointLight((-0.6667 * extent, 50.0, -0.1667 * extent), (0.6, 0.6, 0.6)))
    b.add_light(S.DirectionalLight((-0.3, -1.0, -0.2), (0.3, 0.3, 0.3)))
    b.camera = camera or S.SimplePerspectiveCamera.new((0.0, 14.0, extent * 7.0 / 6.0), (0.0, -0.18, -1.0), (0.0, 1.0, 0.0), 2.0)
    b.background = background or S.SolidColorBackground((0.2, 0.3, 0.5))
    b.options = S.Options(width, height, antialias)
    return b.build()


def spheres_10k(width=3840, height=2160, n=10_000):
    """BASELINE config 3: 10k spheres + ground plane, 4K, depth 8 (pass max_depth=8, seed=3 when rendering)."""
    return random_spheres(n, 0x5EED0003, 60.0, width, height)


def spheres_1m(width=7680, height=4320, n=1_000_000):
    """BASELINE config 4: 1M spheres, 8K, depth 8, device BVH."""
    cam = S.SimplePerspectiveCamera.new((0.0, 40.0, 700.0), (0.0, -0.12, -1.0), (0.0, 1.0, 0.0), 2.0)
    return random_spheres(n, 0x5EED0004, 600.0, width, height, camera=cam)


def _checker_face(size, cell, c0, c1):
    yy, xx = np.mgrid[0:size, 0:size]
    mask = (((xx // cell) + (yy // cell)) & 1).astype(bool)
    out = np.empty((size, size, 3), dtype=np.uint8)
    out[~mask] = c0
    out[mask] = c1
    return out


def _noise_face(size, seed):
    rng = SplitMix64(seed)
    g = 17
    lat = rng.uniform(g * g * 3).reshape(g, g, 3)
    t = np.linspace(0, g - 1, size, endpoint=False)
    i0 = np.floor(t).astype(int)
    f = t - i0
    i1 = np.minimum(i0 + 1, g - 1)
    rows = lat[i0] * (1 - f)[:, None, None] + lat[i1] * f[:, None, None]
    img = rows[:, i0] * (1 - f)[None, :, None] + rows[:, i1] * f[None, :, None]
    grad = np.linspace(0.25, 1.0, size)[:, None, None]
    return np.clip(img * grad * 255.0, 0, 255).astype(np.uint8)


def skybox_faces(size=1024, cell=64):
    """Six synthetic RGB8 faces (px nx py ny pz nz): four checker faces, two value-noise 'image' faces."""
    return [
        _checker_face(size, cell, (230, 230, 230), (40, 60, 160)),
        _checker_face(size, cell, (230, 200, 60), (60, 30, 30)),
        _noise_face(size, 0x5EED0051),
        _checker_face(size, cell, (60, 160, 60), (20, 20, 20)),
        _checker_face(size, cell, (200, 80, 80), (240, 240, 240)),
        _noise_face(size, 0x5EED0052),
    ]


def skybox_100k(width=3840, height=2160, antialias=16, n=100_000, face_size=1024):
    """BASELINE config 5: 100k reflective spheres under a textured skybox (the only texture path upstream), 16 spp."""
    bg = S.SkyboxBackground(*skybox_faces(face_size))
    cam = S.SimplePerspectiveCamera.new((0.0, 20.0, 221.0), (0.0, -0.15, -1.0), (0.0, 1.0, 0.0), 2.0)
    return random_spheres(n, 0x5EED0005, 190.0, width, height, antialias=antialias, camera=cam, reflective_only=True, background=bg)


def materials_showcase(width=256, height=192, antialias=2):
    """Small scene touching every material, light, shape and both inside/outside sphere hits (parity tests)."""
    b = S.SceneBuilder()
    b.add_object(S.Plane((0, 0, 0), (0, 1, 0)), S.PhongMaterial((0.7, 0.7, 0.6), (0.25, 0.25, 0.25), 20.0, (0.02, 0.02, 0.02)))
    b.add_object(S.Plane((0, 0, -9), (0, 0, 1)), S.FresnelMaterial((0.2, 0.3, 0.6), (0.8, 0.8, 0.8), 80.0, (0.01, 0.01, 0.02), 1.8))
    b.add_object(S.Sphere((-2.6, 1.0, -1.0), 1.0), S.PhongMaterial((0.8, 0.2, 0.2), (0.5, 0.5, 0.5), 60.0, (0.02, 0.0, 0.0)))
    b.add_object(S.Sphere((0.0, 1.2, 0.5), 1.2), S.TransparentMaterial((0.9, 0.9, 0.9), 150.0, 1.5))
    b.add_object(S.Sphere((2.7, 0.9, -0.5), 0.9), S.FresnelMaterial((0.1, 0.6, 0.2), (0.9, 0.9, 0.9), 120.0, (0.0, 0.02, 0.0), 1.45))
    b.add_object(S.Sphere((0.8, 0.5, 2.6), 0.5), S.IndirectPhongMaterial((0.9, 0.8, 0.3), (0.0, 0.0, 0.0), 1.0, (0.01, 0.01, 0.0), 3))
    b.add_object(S.Sphere((-1.2, 0.4, 2.2), 0.4), S.TransparentMaterial((0.7, 0.8, 0.9), 90.0, 1.33))
    b.add_light(S.PointLight((4.0, 8.0, 6.0), (0.9, 0.9, 0.9)))
    b.add_light(S.DirectionalLight((0.4, -1.0, -0.3), (0.35, 0.35, 0.4)))
    b.add_light(S.AreaLight((-5.0, 6.0, 2.0), (1.0, 0.0, 0.0), (0.0, 0.0, 1.0), (0.5, 0.45, 0.4)))
    b.camera = S.SimplePerspectiveCamera.look_at((0.0, 1.0, 0.0), (0.0, -0.25, -1.0), (0.0, 1.0, 0.0), np.deg2rad(50.0), 3.0)
    b.background = S.SolidColorBackground((0.15, 0.2, 0.3))
    b.options = S.Options(width, height, antialias)
    return b.build()
