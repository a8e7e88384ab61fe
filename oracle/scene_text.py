"""Pure-Python restatement of the reference's scene-text reader -- TEST INFRASTRUCTURE ONLY.

Follows src/serialize.rs of the upstream crate:
  Location / LL1 (1-token peek, row:col bookkeeping) ..... :22-114
  Acceptor ................................................ :116-172
  expect() and its messages ............................... :174-198
  string escapes .......................................... :295-360
  Tokenizer ............................................... :362-424
  deserialize (lexer error overrides parser error) ........ :427-441
  scalar parsers, warnings ................................ :443-488
  vec3/pnt3/color ......................................... :490-522
  struct / box / function grammars ........................ :524-814
  error Display ........................................... :244-256, :268-272

`deserialize(text)` returns a scene-description dict (schema: rust-raytrace_b200/scene.py) or raises
SyntaxError_ whose str() is the reference's `row:col: message`.  It is the checker for the product's C++
parser (rust-raytrace_b200/csrc/serialize.cpp).

Known, deliberate deviation: an unterminated `/* ...` comment makes the reference loop forever
(src/serialize.rs:395-399); here it raises "invalid token" at end of input.
"""
import math
import re

import numpy as np


class SyntaxError_(Exception):
    def __init__(self, row, col, msg):
        super().__init__("%d:%d: %s" % (row, col, msg))
        self.row, self.col, self.msg = row, col, msg


# ------------------------------------------------------------------------------------------------
# char-level LL1 with location tracking (src/serialize.rs:22-114)
# ------------------------------------------------------------------------------------------------
class _Chars:
    def __init__(self, text):
        self.text = text
        self.i = 0
        self.row, self.col = 1, 0
        self.peeked = None
        self.has_peek = False

    def _pull(self):
        if self.i >= len(self.text):
            return None
        c = self.text[self.i]
        self.i += 1
        if c == "\n":
            self.row += 1
            self.col = 0
        else:
            self.col += 1
        return c

    def peek(self):
        if not self.has_peek:
            c = self._pull()
            if c is None:
                return None  # upstream re-polls the exhausted iterator; same effect
            self.peeked, self.has_peek = c, True
        return self.peeked

    def take(self):
        if self.has_peek:
            self.has_peek = False
            return self.peeked
        return self._pull()

    def accept(self, fn):
        c = self.peek()
        if c is not None and fn(c):
            return self.take()
        return None

    def skip_while(self, fn):
        while True:
            c = self.peek()
            if c is None or not fn(c):
                return
            self.take()

    def take_while(self, fn):
        out = []
        while True:
            c = self.peek()
            if c is None or not fn(c):
                return "".join(out)
            out.append(self.take())


def _is_ws(c):
    return c.isspace()


def _hexdigit(c):
    if c is None:
        return None
    try:
        return int(c, 16) if len(c) == 1 and c in "0123456789abcdefABCDEF" else None
    except ValueError:
        return None


def _parse_string(a):
    """src/serialize.rs:299-355; called after the opening quote has been skipped."""
    out = []
    while True:
        if a.accept(lambda c: c == "\\") is not None:
            c = a.take()
            if c is None:
                break
            simple = {"n": "\n", "r": "\r", "t": "\t", "\\": "\\", "0": "\0", "'": "'", '"': '"'}
            if c in simple:
                out.append(simple[c])
            elif c == "x":
                hi = _hexdigit(a.take())
                if hi is None:
                    a.take()
                    continue
                lo = _hexdigit(a.take())
                if lo is None:
                    continue
                out.append(chr(hi * 16 + lo))
            elif c == "u":
                if a.accept(lambda ch: ch == "{") is None:
                    continue
                acc = 0
                bad = False
                while a.accept(lambda ch: ch == "}") is None:
                    d = _hexdigit(a.take())
                    if d is None:
                        a.skip_while(lambda ch: ch != "}")
                        bad = True
                        break
                    acc = (acc * 16 + d) & 0xFFFFFFFF
                if bad:
                    continue
                if acc < 0x110000 and not (0xD800 <= acc <= 0xDFFF):
                    out.append(chr(acc))
            elif c == "\n":
                a.skip_while(_is_ws)
            # any other escape is dropped
        elif a.accept(lambda c: c == '"') is not None:
            break
        else:
            c = a.take()
            if c is None:
                break
            out.append(c)
    return "".join(out)


_NUM_RE = re.compile(r"^[+-]?(?:(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?|inf|infinity|nan)$", re.IGNORECASE)
_IDENT_START = set("ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz_")
_IDENT_BODY = _IDENT_START | set("0123456789")
_NUM_START = set("0123456789.-+")
_NUM_BODY = _IDENT_BODY | set(".-+")
_PUNCT = {"{": "LBrace", "}": "RBrace", "[": "LBracket", "]": "RBracket", "(": "LParen", ")": "RParen", ":": "Colon", ",": "Comma"}


def _rust_f64_from_str(s):
    if not _NUM_RE.match(s):
        return None
    try:
        return float(s)
    except ValueError:
        return None


class _Tokenizer:
    """src/serialize.rs:362-424.  Tokens are (kind, value) tuples."""

    def __init__(self, text):
        self.a = _Chars(text)
        self.error = None

    def _fail(self, msg):
        self.error = SyntaxError_(self.a.row, self.a.col, msg)
        return None

    def next(self):
        while True:
            if self.error is not None:
                return None
            a = self.a
            a.skip_while(_is_ws)
            c = a.peek()
            if c is None:
                return None
            if c in _PUNCT:
                a.take()
                return (_PUNCT[c], None)
            if c == "#":
                a.skip_while(lambda ch: ch != "\n")
                continue
            if c == "/":
                a.take()
                d = a.take()
                if d == "/":
                    a.skip_while(lambda ch: ch != "\n")
                elif d == "*":
                    while True:
                        a.skip_while(lambda ch: ch != "*")
                        a.take()
                        e = a.take()
                        if e == "/":
                            break
                        if e is None and a.peek() is None:
                            return self._fail("invalid token")  # upstream would spin forever here
                else:
                    return self._fail("invalid token")
                continue
            if c == '"':
                a.take()
                return ("String", _parse_string(a))
            if c in _IDENT_START:
                return ("Identifier", a.take_while(lambda ch: ch in _IDENT_BODY))
            if c in _NUM_START:
                num = a.take_while(lambda ch: ch in _NUM_BODY)
                v = _rust_f64_from_str(num)
                if v is None:
                    return self._fail("invalid number: %s" % num)
                return ("Number", v)
            return self._fail("invalid token")


def _fmt_f64_debug(x):
    """Rust `{:?}` for f64 (shortest round-trip; exponent form outside [1e-5, 1e16))."""
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "inf" if x > 0 else "-inf"
    ax = abs(x)
    if ax != 0.0 and (ax < 1e-5 or ax >= 1e16):
        m, e = ("%r" % x).lower().split("e") if "e" in ("%r" % x).lower() else (None, None)
        if m is None:
            s = "%e" % x
            m, e = s.split("e")
            m = repr(float(m))
        if m.endswith(".0"):
            m = m[:-2]
        return "%se%d" % (m, int(e))
    r = repr(float(x))
    if "e" in r or "E" in r:
        r = ("%f" % x).rstrip("0")
        if r.endswith("."):
            r += "0"
    return r


def _tok_debug(tok):
    kind, v = tok
    if kind == "Identifier" or kind == "String":
        esc = v.replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n").replace("\t", "\\t").replace("\r", "\\r")
        return '%s("%s")' % (kind, esc)
    if kind == "Number":
        return "Number(%s)" % _fmt_f64_debug(v)
    return kind


class _Toks:
    """Token-level LL1 + Acceptor + expect (src/serialize.rs:67-198, 419-424)."""

    def __init__(self, text):
        self.tz = _Tokenizer(text)
        self.row, self.col = 1, 0
        self.peeked = None
        self.has_peek = False
        self.warnings = []

    def _pull(self):
        t = self.tz.next()
        self.row, self.col = self.tz.a.row, self.tz.a.col
        return t

    def peek(self):
        if not self.has_peek:
            t = self._pull()
            if t is None:
                return None
            self.peeked, self.has_peek = t, True
        return self.peeked

    def take(self):
        if self.has_peek:
            self.has_peek = False
            return self.peeked
        return self._pull()

    def accept(self, kind):
        t = self.peek()
        if t is not None and t[0] == kind:
            return self.take()
        return None

    def err(self, msg):
        return SyntaxError_(self.row, self.col, msg)

    def expect(self, fn, desc):
        t = self.peek()
        if t is None:
            raise self.err("expected %s (end of file)" % desc)
        if fn(t):
            return self.take()
        raise self.err("expected %s, not %s" % (desc, _tok_debug(t)))

    def expect_kind(self, kind):
        return self.expect(lambda t: t[0] == kind, kind)


# ------------------------------------------------------------------------------------------------
# scalar parsers (src/serialize.rs:443-522)
# ------------------------------------------------------------------------------------------------
def _f64(t):
    return t.expect_kind("Number")[1]


def _rust_round(x):
    if x != x or x in (float("inf"), float("-inf")):
        return x
    return math.floor(x + 0.5) if x >= 0 else -math.floor(-x + 0.5)


def _i32(t):
    num = _f64(t)
    frac = abs(math.fmod(num, 1.0)) if num == num and abs(num) != float("inf") else float("nan")
    if frac > 0.01:
        t.warnings.append("Warning: %s stored as integer" % _fmt_display(num))
    if abs(num) > 1677215.0:
        t.warnings.append("Warning: integer values past ~2^24+1 are not exact")
    r = _rust_round(num)
    if r != r:
        return 0
    return int(max(-2147483648, min(2147483647, r)))


def _fmt_display(x):
    r = repr(float(x))
    return r[:-2] if r.endswith(".0") else r


def _u32(t):
    num = _i32(t)
    if num < 0:
        t.warnings.append("Warning: unsigned integer %d is negative, using 0" % num)
        return 0
    return num


def _string(t):
    return t.expect_kind("String")[1]


def _ang(t):
    num = _f64(t)
    unit = t.expect_kind("Identifier")[1]
    if unit == "deg":
        return num * math.pi / 180.0
    if unit == "rad":
        return num
    raise t.err("no such class: %s" % unit)


def _vec3(t):
    t.expect_kind("LParen")
    x = _f64(t)
    t.expect_kind("Comma")
    y = _f64(t)
    t.expect_kind("Comma")
    z = _f64(t)
    t.expect_kind("RParen")
    return (x, y, z)


def _color(t):
    t.expect(lambda k: k[0] == "Identifier" and k[1] == "rgb", 'Identifier("rgb")')
    return _vec3(t)


# ------------------------------------------------------------------------------------------------
# grammar macros (src/serialize.rs:524-604)
# ------------------------------------------------------------------------------------------------
def _struct(t, fields):
    """fn_parse_struct!: `{ name: value ... }`, any order, repeats allowed (last wins), all required."""
    t.expect_kind("LBrace")
    got = {}
    while t.accept("RBrace") is None:
        name = t.expect_kind("Identifier")[1]
        if name not in fields:
            raise t.err("undefined field: %s" % name)
        t.expect_kind("Colon")
        got[name] = fields[name](t)
    if len(got) != len(fields):
        raise t.err("missing one or more fields")
    return got


def _box(t, classes):
    cls = t.expect_kind("Identifier")[1]
    if cls not in classes:
        raise t.err("no such class: %s" % cls)
    return classes[cls](t)


def _function(t, fname, params):
    t.expect(lambda k: k[0] == "Identifier" and k[1] == fname, 'Identifier("%s")' % fname)
    t.expect_kind("LParen")
    out = []
    for i, p in enumerate(params):
        if i:
            t.expect_kind("Comma")
        out.append(p(t))
    t.expect_kind("RParen")
    return out


def _vec(t, parser):
    t.expect_kind("LBracket")
    out = []
    while t.accept("RBracket") is None:
        out.append(parser(t))
    return out


# ------------------------------------------------------------------------------------------------
# per-type grammars (src/serialize.rs:606-814)
# ------------------------------------------------------------------------------------------------
def _shape(t):
    def sphere(t):
        d = _struct(t, {"center": _vec3, "radius": _f64})
        return {"shape": "sphere", "center": d["center"], "radius": d["radius"]}

    def plane(t):
        d = _struct(t, {"point": _vec3, "normal": _vec3})
        return {"shape": "plane", "point": d["point"], "normal": d["normal"]}

    return _box(t, {"Sphere": sphere, "Plane": plane})


def _new_spc(t):
    p = _function(t, "new", [_vec3, _vec3, _vec3, _f64])
    return {"ctor": "new", "position": p[0], "look": p[1], "up": p[2], "im_dist": p[3]}


def _look_at_spc(t):
    p = _function(t, "look_at", [_vec3, _vec3, _vec3, _ang, _f64])
    return {"ctor": "look_at", "focus": p[0], "look": p[1], "up": p[2], "pov": p[3], "h": p[4]}


def _spc(t):
    # parse_new_spc(toks).or_else(|_| parse_look_at_spc(toks)) -- no backtracking
    try:
        return _new_spc(t)
    except SyntaxError_:
        return _look_at_spc(t)


def _dofc(t):
    p = _function(t, "new", [_spc, _f64, _f64, _u32])
    cam = dict(p[0])
    cam["dof"] = {"focus": p[1], "aperture": p[2], "samples": p[3]}
    return cam


def _camera(t):
    return _box(t, {"SimplePerspectiveCamera": _spc, "DepthOfFieldCamera": _dofc})


def _material(t):
    def phong(t):
        d = _struct(t, {"diffuse": _color, "specular": _color, "exponent": _f64, "ambient": _color})
        return dict(d, kind="phong")

    def indirect(t):
        d = _struct(t, {"diffuse": _color, "specular": _color, "exponent": _f64, "ambient": _color, "samples": _u32})
        return dict(d, kind="indirect")

    def fresnel(t):
        d = _struct(t, {"diffuse": _color, "specular": _color, "exponent": _f64, "ambient": _color, "ior": _f64})
        return dict(d, kind="fresnel")

    def transparent(t):
        d = _struct(t, {"specular": _color, "exponent": _f64, "ior": _f64})
        return dict(d, kind="transparent")

    return _box(t, {"PhongMaterial": phong, "IndirectPhongMaterial": indirect, "FresnelMaterial": fresnel, "TransparentMaterial": transparent})


def _object(t):
    return _struct(t, {"bounds": _shape, "material": _material})


def _light_model(t):
    def point(t):
        return dict(_struct(t, {"location": _vec3}), model="point")

    def directional(t):
        return dict(_struct(t, {"direction": _vec3}), model="directional")

    def area(t):
        return dict(_struct(t, {"origin": _vec3, "side1": _vec3, "side2": _vec3}), model="area")

    return _box(t, {"PointLight": point, "DirectionalLight": directional, "AreaLight": area})


def _light(t):
    d = _struct(t, {"model": _light_model, "color": _color})
    return dict(d["model"], color=d["color"])


def _make_background_parser(texture_loader):
    def load_texture(t):
        path = _function(t, "load", [_string])[0]
        try:
            return np.ascontiguousarray(texture_loader(path), dtype=np.uint8)
        except Exception as e:  # noqa: BLE001 -- any loader failure is a TextureLoad error upstream
            raise t.err('error loading "%s": %s' % (path, e))

    def solid(t):
        return {"type": "solid", "color": _struct(t, {"color": _color})["color"]}

    def skybox(t):
        d = _struct(t, {k: load_texture for k in ("px", "nx", "py", "ny", "pz", "nz")})
        return {"type": "skybox", "faces": [d[k] for k in ("px", "nx", "py", "ny", "pz", "nz")]}

    return lambda t: _box(t, {"SolidColorBackground": solid, "SkyboxBackground": skybox})


def _options(t):
    return _struct(t, {"width": _u32, "height": _u32, "antialias": _u32})


def _default_texture_loader(path):
    from PIL import Image

    return np.asarray(Image.open(path).convert("RGB"))


def deserialize(text, texture_loader=None, return_warnings=False):
    """src/serialize.rs:427-441."""
    t = _Toks(text)
    bg = _make_background_parser(texture_loader or _default_texture_loader)
    result, perr = None, None
    try:
        result = _struct(t, {"objects": lambda t: _vec(t, _object), "lights": lambda t: _vec(t, _light), "camera": _camera,
                             "background": bg, "options": _options})
    except SyntaxError_ as e:
        perr = e
    if t.tz.error is not None:
        raise t.tz.error
    if perr is not None:
        raise perr
    materials, objects = [], []
    for o in result["objects"]:
        materials.append(o["material"])
        objects.append(dict(o["bounds"], material=len(materials) - 1))
    desc = {"materials": materials, "objects": objects, "lights": result["lights"], "camera": result["camera"],
            "background": result["background"], "options": result["options"]}
    return (desc, t.warnings) if return_warnings else desc
