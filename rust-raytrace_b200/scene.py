"""Scene descriptions: plain dicts that mirror the reference's `pub` scene types, plus a writer for the text format.

Schema (all vectors/colours are 3-sequences of floats):

    {"materials": [ {"kind": "phong"|"indirect"|"fresnel"|"transparent", "diffuse", "specular", "exponent",
                     "ambient", "samples" (indirect), "ior" (fresnel, transparent)} ... ],
     "objects":   [ {"shape": "sphere", "center", "radius", "material": index}
                  | {"shape": "plane", "point", "normal", "material": index}
                  | {"shape": "spheres", "data": float64 (n,4) centre+radius, "material": uint32 (n,)} ... ],
     "lights":    [ {"model": "point", "location", "color"} | {"model": "directional", "direction", "color"}
                  | {"model": "area", "origin", "side1", "side2", "color"} ... ],
     "camera":    {"ctor": "new", "position", "look", "up", "im_dist"}
                | {"ctor": "look_at", "focus", "look", "up", "pov" (radians), "h"}     [+ "dof": {"focus", "aperture", "samples"}],
     "background": {"type": "solid", "color"} | {"type": "skybox", "faces": [uint8 (h,w,3)] * 6  (px nx py ny pz nz)},
     "options":   {"width", "height", "antialias"}}

Object order is `Scene.objects` order upstream (src/scene.rs:203): it is the id reported for primary hits and the
tie-break of the closest-hit search.  The helper functions below carry the upstream type names
(src/shapes.rs:43-98, src/scene.rs:32-198, src/camera.rs:31-106).
"""
import numpy as np


def _v(x):
    return tuple(float(c) for c in x)


# ---- materials (src/scene.rs:32-89) ----------------------------------------------------------------------
def PhongMaterial(diffuse, specular, exponent, ambient):
    return {"kind": "phong", "diffuse": _v(diffuse), "specular": _v(specular), "exponent": float(exponent), "ambient": _v(ambient)}


def IndirectPhongMaterial(diffuse, specular, exponent, ambient, samples):
    return {"kind": "indirect", "diffuse": _v(diffuse), "specular": _v(specular), "exponent": float(exponent), "ambient": _v(ambient),
            "samples": int(samples)}


def FresnelMaterial(diffuse, specular, exponent, ambient, ior):
    return {"kind": "fresnel", "diffuse": _v(diffuse), "specular": _v(specular), "exponent": float(exponent), "ambient": _v(ambient),
            "ior": float(ior)}


def TransparentMaterial(specular, exponent, ior):
    return {"kind": "transparent", "specular": _v(specular), "exponent": float(exponent), "ior": float(ior)}


# ---- shapes (src/shapes.rs:43-48, :93-98) -----------------------------------------------------------------
def Sphere(center, radius):
    return {"shape": "sphere", "center": _v(center), "radius": float(radius)}


def Plane(point, normal):
    return {"shape": "plane", "point": _v(point), "normal": _v(normal)}


# ---- lights (src/scene.rs:109-155) ------------------------------------------------------------------------
def PointLight(location, color):
    return {"model": "point", "location": _v(location), "color": _v(color)}


def DirectionalLight(direction, color):
    return {"model": "directional", "direction": _v(direction), "color": _v(color)}


def AreaLight(origin, side1, side2, color):
    return {"model": "area", "origin": _v(origin), "side1": _v(side1), "side2": _v(side2), "color": _v(color)}


# ---- cameras (src/camera.rs:51-73, :97-106) ---------------------------------------------------------------
class SimplePerspectiveCamera:
    @staticmethod
    def new(position, look, up, im_dist):
        return {"ctor": "new", "position": _v(position), "look": _v(look), "up": _v(up), "im_dist": float(im_dist)}

    @staticmethod
    def look_at(focus, look, up, pov, h):
        return {"ctor": "look_at", "focus": _v(focus), "look": _v(look), "up": _v(up), "pov": float(pov), "h": float(h)}


class DepthOfFieldCamera:
    @staticmethod
    def new(camera, focus, aperture, samples):
        return dict(camera, dof={"focus": float(focus), "aperture": float(aperture), "samples": int(samples)})


# ---- backgrounds / options (src/scene.rs:165-198) ---------------------------------------------------------
def SolidColorBackground(color):
    return {"type": "solid", "color": _v(color)}


def SkyboxBackground(px, nx, py, ny, pz, nz):
    return {"type": "skybox", "faces": [np.ascontiguousarray(f, dtype=np.uint8) for f in (px, nx, py, ny, pz, nz)]}


def Options(width, height, antialias):
    return {"width": int(width), "height": int(height), "antialias": int(antialias)}


class SceneBuilder:
    """Collects objects with their materials the way `Object {bounds, material}` pairs them upstream (src/scene.rs:93-98)."""

    def __init__(self):
        self.materials, self.objects, self.lights = [], [], []
        self.camera = self.background = self.options = None

    def add_material(self, material):
        self.materials.append(material)
        return len(self.materials) - 1

    def add_object(self, bounds, material):
        """material: a material dict (appended) or an index into the materials added so far"""
        idx = material if isinstance(material, int) else self.add_material(material)
        self.objects.append(dict(bounds, material=idx))
        return len(self.objects) - 1

    def add_spheres(self, center_radius, material_index):
        data = np.ascontiguousarray(center_radius, dtype=np.float64).reshape(-1, 4)
        mat = np.ascontiguousarray(material_index, dtype=np.uint32).reshape(-1)
        assert mat.shape[0] == data.shape[0]
        self.objects.append({"shape": "spheres", "data": data, "material": mat})

    def add_light(self, light):
        self.lights.append(light)

    def build(self):
        return {"materials": self.materials, "objects": self.objects, "lights": self.lights, "camera": self.camera,
                "background": self.background, "options": self.options}


def count_objects(desc):
    return sum(o["data"].shape[0] if o["shape"] == "spheres" else 1 for o in desc["objects"])


# ---- text writer: emits the grammar of src/serialize.rs:606-814 --------------------------------------------
def _f(x):
    x = float(x)
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "inf" if x > 0 else "-inf"
    return repr(x)


def _t3(v):
    return "(%s, %s, %s)" % (_f(v[0]), _f(v[1]), _f(v[2]))


def _rgb(v):
    return "rgb" + _t3(v)


def _material_text(m):
    k = m["kind"]
    if k == "phong":
        return "PhongMaterial { diffuse: %s specular: %s exponent: %s ambient: %s }" % (_rgb(m["diffuse"]), _rgb(m["specular"]), _f(m["exponent"]), _rgb(m["ambient"]))
    if k == "indirect":
        return "IndirectPhongMaterial { diffuse: %s specular: %s exponent: %s ambient: %s samples: %d }" % (
            _rgb(m["diffuse"]), _rgb(m["specular"]), _f(m["exponent"]), _rgb(m["ambient"]), m["samples"])
    if k == "fresnel":
        return "FresnelMaterial { diffuse: %s specular: %s exponent: %s ambient: %s ior: %s }" % (
            _rgb(m["diffuse"]), _rgb(m["specular"]), _f(m["exponent"]), _rgb(m["ambient"]), _f(m["ior"]))
    return "TransparentMaterial { specular: %s exponent: %s ior: %s }" % (_rgb(m["specular"]), _f(m["exponent"]), _f(m["ior"]))


def _spc_text(cam):
    if cam["ctor"] == "new":
        return "new(%s, %s, %s, %s)" % (_t3(cam["position"]), _t3(cam["look"]), _t3(cam["up"]), _f(cam["im_dist"]))
    return "look_at(%s, %s, %s, %s rad, %s)" % (_t3(cam["focus"]), _t3(cam["look"]), _t3(cam["up"]), _f(cam["pov"]), _f(cam["h"]))


def to_text(desc, texture_paths=None):
    """Scene text that `serialize::deserialize` reads back into the same scene (numbers via repr: exact round trip)."""
    out = ["{", "  objects: ["]
    mats = desc["materials"]
    for o in desc["objects"]:
        if o["shape"] == "spheres":
            for (cx, cy, cz, r), mi in zip(o["data"], o["material"]):
                out.append("    { bounds: Sphere { center: %s radius: %s } material: %s }" % (_t3((cx, cy, cz)), _f(r), _material_text(mats[int(mi)])))
        elif o["shape"] == "sphere":
            out.append("    { bounds: Sphere { center: %s radius: %s } material: %s }" % (_t3(o["center"]), _f(o["radius"]), _material_text(mats[o["material"]])))
        else:
            out.append("    { bounds: Plane { point: %s normal: %s } material: %s }" % (_t3(o["point"]), _t3(o["normal"]), _material_text(mats[o["material"]])))
    out.append("  ]")
    out.append("  lights: [")
    for l in desc["lights"]:
        if l["model"] == "point":
            m = "PointLight { location: %s }" % _t3(l["location"])
        elif l["model"] == "directional":
            m = "DirectionalLight { direction: %s }" % _t3(l["direction"])
        else:
            m = "AreaLight { origin: %s side1: %s side2: %s }" % (_t3(l["origin"]), _t3(l["side1"]), _t3(l["side2"]))
        out.append("    { model: %s color: %s }" % (m, _rgb(l["color"])))
    out.append("  ]")
    cam = desc["camera"]
    if cam.get("dof"):
        d = cam["dof"]
        out.append("  camera: DepthOfFieldCamera new(%s, %s, %s, %d)" % (_spc_text(cam), _f(d["focus"]), _f(d["aperture"]), d["samples"]))
    else:
        out.append("  camera: SimplePerspectiveCamera %s" % _spc_text(cam))
    bg = desc["background"]
    if bg["type"] == "solid":
        out.append("  background: SolidColorBackground { color: %s }" % _rgb(bg["color"]))
    else:
        if not texture_paths or len(texture_paths) != 6:
            raise ValueError("a skybox scene needs six texture paths to be written as text")
        names = ("px", "nx", "py", "ny", "pz", "nz")
        out.append("  background: SkyboxBackground { %s }" % " ".join('%s: load("%s")' % (n, p) for n, p in zip(names, texture_paths)))
    o = desc["options"]
    out.append("  options: { width: %d height: %d antialias: %d }" % (o["width"], o["height"], o["antialias"]))
    out.append("}")
    return "\n".join(out) + "\n"
