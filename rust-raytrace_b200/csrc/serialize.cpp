// serialize.cpp -- scene text reader: the host-side mirror of `serialize::deserialize` (src/serialize.rs:427-441).
//
// Same token set, grammar, location bookkeeping and error strings as upstream, so that a scene file and its
// error messages behave identically (`row:col: message`, src/serialize.rs:244-272):
//   - locations: 1-based row, column of the last character pulled from the input (including the one-character
//     lookahead of the lexer), src/serialize.rs:22-114;
//   - lexer errors take priority over parser errors, src/serialize.rs:436-440;
//   - struct fields in any order, repeats allowed (last wins), all required, src/serialize.rs:524-553;
//   - `new(..)` then `look_at(..)` tried in turn for cameras without backtracking, src/serialize.rs:648-663.
// Runs once per scene on the host; not part of the accelerated path.
//
// Deliberate deviation: an unterminated `/*` comment spins forever upstream (src/serialize.rs:395-399);
// here it is reported as "invalid token".
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "../../include/raytrace_b200.h"
#include "host_scene.h"

namespace rt {
namespace {

struct ParseError {
    size_t row, col;
    std::string msg;
};

// decode one UTF-8 scalar starting at s[i]; returns its length (>= 1) and the code point
size_t utf8_decode(const char* s, size_t len, size_t i, uint32_t* cp) {
    unsigned char c = (unsigned char)s[i];
    if (c < 0x80) { *cp = c; return 1; }
    int extra = (c >= 0xF0) ? 3 : (c >= 0xE0) ? 2 : (c >= 0xC0) ? 1 : 0;
    uint32_t v = extra == 3 ? (c & 0x07) : extra == 2 ? (c & 0x0F) : extra == 1 ? (c & 0x1F) : c;
    size_t n = 1;
    for (int k = 0; k < extra && i + n < len; k++, n++) v = (v << 6) | ((unsigned char)s[i + n] & 0x3F);
    *cp = v;
    return n;
}
void utf8_encode(uint32_t cp, std::string* out) {
    if (cp < 0x80) out->push_back((char)cp);
    else if (cp < 0x800) { out->push_back((char)(0xC0 | (cp >> 6))); out->push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) { out->push_back((char)(0xE0 | (cp >> 12))); out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F))); }
    else { out->push_back((char)(0xF0 | (cp >> 18))); out->push_back((char)(0x80 | ((cp >> 12) & 0x3F))); out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F))); }
}
bool is_whitespace(uint32_t c) {  // char::is_whitespace
    return c == ' ' || (c >= 0x09 && c <= 0x0D) || c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 ||
           c == 0x202F || c == 0x205F || c == 0x3000;
}

constexpr uint32_t END = 0xFFFFFFFFu;

// character stream with one-character peek and upstream's location rule
class CharStream {
public:
    CharStream(const char* s, size_t len) : s_(s), len_(len) {}
    size_t row = 1, col = 0;
    uint32_t peek() {
        if (!has_peek_) {
            uint32_t c = pull();
            if (c == END) return END;
            peeked_ = c;
            has_peek_ = true;
        }
        return peeked_;
    }
    uint32_t take() {
        if (has_peek_) { has_peek_ = false; return peeked_; }
        return pull();
    }
    template <typename F>
    bool accept(F f) {
        uint32_t c = peek();
        if (c != END && f(c)) { take(); return true; }
        return false;
    }
    template <typename F>
    void skip_while(F f) {
        for (;;) {
            uint32_t c = peek();
            if (c == END || !f(c)) return;
            take();
        }
    }
    template <typename F>
    std::string take_while(F f) {
        std::string out;
        for (;;) {
            uint32_t c = peek();
            if (c == END || !f(c)) return out;
            utf8_encode(take(), &out);
        }
    }

private:
    uint32_t pull() {
        if (pos_ >= len_) return END;
        uint32_t cp;
        pos_ += utf8_decode(s_, len_, pos_, &cp);
        if (cp == '\n') { row++; col = 0; } else { col++; }
        return cp;
    }
    const char* s_;
    size_t len_, pos_ = 0;
    uint32_t peeked_ = 0;
    bool has_peek_ = false;
};

enum class Tok { Identifier, String, Number, LBrace, RBrace, LBracket, RBracket, LParen, RParen, Colon, Comma, End };
const char* tok_name(Tok t) {
    static const char* N[] = {"Identifier", "String", "Number", "LBrace", "RBrace", "LBracket", "RBracket", "LParen", "RParen", "Colon", "Comma", "End"};
    return N[(int)t];
}
struct Token {
    Tok kind = Tok::End;
    std::string text;
    double number = 0.0;
};

int hex_digit(uint32_t c) {
    if (c >= '0' && c <= '9') return (int)(c - '0');
    if (c >= 'a' && c <= 'f') return (int)(c - 'a' + 10);
    if (c >= 'A' && c <= 'F') return (int)(c - 'A' + 10);
    return -1;
}

// src/serialize.rs:299-355 (opening quote already consumed)
std::string lex_string(CharStream& a) {
    std::string out;
    for (;;) {
        if (a.accept([](uint32_t c) { return c == '\\'; })) {
            uint32_t c = a.take();
            if (c == END) break;
            switch (c) {
            case 'n': out.push_back('\n'); break;
            case 'r': out.push_back('\r'); break;
            case 't': out.push_back('\t'); break;
            case '\\': out.push_back('\\'); break;
            case '0': out.push_back('\0'); break;
            case '\'': out.push_back('\''); break;
            case '"': out.push_back('"'); break;
            case 'x': {
                uint32_t h = a.take();
                int hi = h == END ? -1 : hex_digit(h);
                if (hi < 0) { a.take(); break; }
                uint32_t l = a.take();
                int lo = l == END ? -1 : hex_digit(l);
                if (lo < 0) break;
                utf8_encode((uint32_t)(hi * 16 + lo), &out);
                break;
            }
            case 'u': {
                if (!a.accept([](uint32_t ch) { return ch == '{'; })) break;
                uint32_t acc = 0;
                bool bad = false;
                while (!a.accept([](uint32_t ch) { return ch == '}'; })) {
                    uint32_t dch = a.take();
                    int d = dch == END ? -1 : hex_digit(dch);
                    if (d < 0) { a.skip_while([](uint32_t ch) { return ch != '}'; }); bad = true; break; }
                    acc = acc * 16 + (uint32_t)d;
                }
                if (!bad && acc < 0x110000 && !(acc >= 0xD800 && acc <= 0xDFFF)) utf8_encode(acc, &out);
                break;
            }
            case '\n': a.skip_while(is_whitespace); break;
            default: break;  // unknown escapes are dropped
            }
        } else if (a.accept([](uint32_t c) { return c == '"'; })) {
            break;
        } else {
            uint32_t c = a.take();
            if (c == END) break;
            utf8_encode(c, &out);
        }
    }
    return out;
}

bool is_ident_start(uint32_t c) { return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z') || c == '_'; }
bool is_ident_body(uint32_t c) { return is_ident_start(c) || (c >= '0' && c <= '9'); }
bool is_num_start(uint32_t c) { return (c >= '0' && c <= '9') || c == '.' || c == '-' || c == '+'; }
bool is_num_body(uint32_t c) { return is_ident_body(c) || c == '.' || c == '-' || c == '+'; }

// what Rust's f64::from_str accepts: [+-]? (digits [. digits?]? | . digits) ([eE][+-]?digits)? | inf | infinity | nan
bool rust_parse_f64(const std::string& s, double* out) {
    size_t i = 0, n = s.size();
    if (i < n && (s[i] == '+' || s[i] == '-')) i++;
    std::string rest = s.substr(i);
    std::string lower;
    for (char c : rest) lower.push_back((char)std::tolower((unsigned char)c));
    if (lower == "inf" || lower == "infinity") { *out = (s[0] == '-') ? -INFINITY : INFINITY; return true; }
    if (lower == "nan") { *out = NAN; return true; }
    size_t j = i, digits = 0;
    while (j < n && s[j] >= '0' && s[j] <= '9') { j++; digits++; }
    if (j < n && s[j] == '.') {
        j++;
        while (j < n && s[j] >= '0' && s[j] <= '9') { j++; digits++; }
    }
    if (digits == 0) return false;
    if (j < n && (s[j] == 'e' || s[j] == 'E')) {
        j++;
        if (j < n && (s[j] == '+' || s[j] == '-')) j++;
        size_t ed = 0;
        while (j < n && s[j] >= '0' && s[j] <= '9') { j++; ed++; }
        if (ed == 0) return false;
    }
    if (j != n) return false;
    *out = std::strtod(s.c_str(), nullptr);
    return true;
}

// src/serialize.rs:362-424
class Lexer {
public:
    Lexer(const char* s, size_t len) : a(s, len) {}
    CharStream a;
    bool failed = false;
    ParseError error{};

    bool next(Token* out) {
        for (;;) {
            if (failed) return false;
            a.skip_while(is_whitespace);
            uint32_t c = a.peek();
            if (c == END) return false;
            Tok single = Tok::End;
            switch (c) {
            case '{': single = Tok::LBrace; break;
            case '}': single = Tok::RBrace; break;
            case '[': single = Tok::LBracket; break;
            case ']': single = Tok::RBracket; break;
            case '(': single = Tok::LParen; break;
            case ')': single = Tok::RParen; break;
            case ':': single = Tok::Colon; break;
            case ',': single = Tok::Comma; break;
            default: break;
            }
            if (single != Tok::End) { a.take(); out->kind = single; return true; }
            if (c == '#') { a.skip_while([](uint32_t ch) { return ch != '\n'; }); continue; }
            if (c == '/') {
                a.take();
                uint32_t d = a.take();
                if (d == '/') {
                    a.skip_while([](uint32_t ch) { return ch != '\n'; });
                } else if (d == '*') {
                    for (;;) {
                        a.skip_while([](uint32_t ch) { return ch != '*'; });
                        a.take();
                        uint32_t e = a.take();
                        if (e == '/') break;
                        if (e == END && a.peek() == END) return fail("invalid token");
                    }
                } else {
                    return fail("invalid token");
                }
                continue;
            }
            if (c == '"') { a.take(); out->kind = Tok::String; out->text = lex_string(a); return true; }
            if (is_ident_start(c)) { out->kind = Tok::Identifier; out->text = a.take_while(is_ident_body); return true; }
            if (is_num_start(c)) {
                std::string num = a.take_while(is_num_body);
                double v;
                if (!rust_parse_f64(num, &v)) return fail("invalid number: " + num);
                out->kind = Tok::Number; out->number = v;
                return true;
            }
            return fail("invalid token");
        }
    }

private:
    bool fail(const std::string& msg) {
        failed = true;
        error = ParseError{a.row, a.col, msg};
        return false;
    }
};

// Rust `{:?}` of an f64
std::string debug_f64(double x) {
    if (x != x) return "NaN";
    if (std::isinf(x)) return x > 0 ? "inf" : "-inf";
    char buf[64];
    for (int prec = 1; prec <= 17; prec++) {
        std::snprintf(buf, sizeof buf, "%.*e", prec - 1, x);
        if (std::strtod(buf, nullptr) == x) break;
    }
    // buf = d.ddddde[+-]XX : split mantissa digits and exponent
    std::string s(buf);
    size_t epos = s.find('e');
    std::string mant = s.substr(0, epos);
    int exp10 = std::atoi(s.c_str() + epos + 1);
    bool neg = mant[0] == '-';
    if (neg) mant.erase(0, 1);
    std::string digits;
    for (char c : mant) if (c != '.') digits.push_back(c);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = neg ? "-" : "";
    const double ax = std::fabs(x);
    if (ax != 0.0 && (ax < 1e-5 || ax >= 1e16)) {
        out += digits.substr(0, 1);
        if (digits.size() > 1) out += "." + digits.substr(1);
        out += "e" + std::to_string(exp10);
        return out;
    }
    if (x == 0.0) return out + "0.0";
    if (exp10 >= 0) {
        std::string ip = digits.substr(0, std::min(digits.size(), (size_t)exp10 + 1));
        while (ip.size() < (size_t)exp10 + 1) ip.push_back('0');
        std::string fp = digits.size() > (size_t)exp10 + 1 ? digits.substr(exp10 + 1) : "0";
        return out + ip + "." + fp;
    }
    return out + "0." + std::string((size_t)(-exp10 - 1), '0') + digits;
}

std::string display_f64(double x) {
    std::string s = debug_f64(x);
    if (s.size() > 2 && s.compare(s.size() - 2, 2, ".0") == 0) s.erase(s.size() - 2);
    return s;
}

std::string debug_str(const std::string& v) {
    std::string o = "\"";
    for (char c : v) {
        switch (c) {
        case '\\': o += "\\\\"; break;
        case '"': o += "\\\""; break;
        case '\n': o += "\\n"; break;
        case '\t': o += "\\t"; break;
        case '\r': o += "\\r"; break;
        default: o.push_back(c);
        }
    }
    return o + "\"";
}

std::string debug_token(const Token& t) {
    switch (t.kind) {
    case Tok::Identifier: return "Identifier(" + debug_str(t.text) + ")";
    case Tok::String: return "String(" + debug_str(t.text) + ")";
    case Tok::Number: return "Number(" + debug_f64(t.number) + ")";
    default: return tok_name(t.kind);
    }
}

struct V3d {
    double v[3];
};

// token stream with one-token peek + the recursive-descent grammar
class Parser {
public:
    Parser(const char* s, size_t len, rt_scene* scene) : lex(s, len), sc(scene) {}
    Lexer lex;
    rt_scene* sc;
    size_t row = 1, col = 0;

    Token* peek() {
        if (!has_peek_) {
            Token t;
            bool ok = pull(&t);
            if (!ok) return nullptr;
            peeked_ = t;
            has_peek_ = true;
        }
        return &peeked_;
    }
    bool take(Token* out) {
        if (has_peek_) { has_peek_ = false; *out = peeked_; return true; }
        return pull(out);
    }
    bool accept(Tok k) {
        Token* t = peek();
        if (t && t->kind == k) { has_peek_ = false; return true; }
        return false;
    }
    [[noreturn]] void fail(const std::string& msg) { throw ParseError{row, col, msg}; }

    Token expect(const std::function<bool(const Token&)>& f, const std::string& desc) {
        Token* t = peek();
        if (!t) fail("expected " + desc + " (end of file)");
        if (!f(*t)) fail("expected " + desc + ", not " + debug_token(*t));
        Token out;
        take(&out);
        return out;
    }
    Token expect_kind(Tok k) {
        return expect([k](const Token& t) { return t.kind == k; }, tok_name(k));
    }
    Token expect_ident(const char* name) {
        std::string n(name);
        return expect([n](const Token& t) { return t.kind == Tok::Identifier && t.text == n; }, "Identifier(\"" + n + "\")");
    }

    // scalars, src/serialize.rs:443-488
    double f64() { return expect_kind(Tok::Number).number; }
    int32_t i32() {
        double num = f64();
        double frac = std::fabs(num - std::trunc(num));
        if (frac > 0.01) sc->warnings += "Warning: " + display_f64(num) + " stored as integer\n";
        if (std::fabs(num) > 1677215.0) sc->warnings += "Warning: integer values past ~2^24+1 are not exact\n";
        double r = std::round(num);  // half away from zero, like f64::round
        if (r != r) return 0;
        if (r >= 2147483647.0) return 2147483647;
        if (r <= -2147483648.0) return (-2147483647 - 1);
        return (int32_t)r;
    }
    uint32_t u32() {
        int32_t n = i32();
        if (n < 0) { sc->warnings += "Warning: unsigned integer " + std::to_string(n) + " is negative, using 0\n"; return 0; }
        return (uint32_t)n;
    }
    std::string str() { return expect_kind(Tok::String).text; }
    double ang() {
        double num = f64();
        std::string unit = expect_kind(Tok::Identifier).text;
        if (unit == "deg") return num * 3.14159265358979323846264338327950288 / 180.0;
        if (unit == "rad") return num;
        fail("no such class: " + unit);
    }
    V3d vec3() {
        V3d r;
        expect_kind(Tok::LParen);
        r.v[0] = f64(); expect_kind(Tok::Comma);
        r.v[1] = f64(); expect_kind(Tok::Comma);
        r.v[2] = f64();
        expect_kind(Tok::RParen);
        return r;
    }
    V3d color() { expect_ident("rgb"); return vec3(); }

    // fn_parse_struct!, src/serialize.rs:524-553
    using FieldFn = std::function<void()>;
    void parse_struct(const std::vector<std::pair<const char*, FieldFn>>& fields) {
        expect_kind(Tok::LBrace);
        std::vector<bool> seen(fields.size(), false);
        while (!accept(Tok::RBrace)) {
            std::string name = expect_kind(Tok::Identifier).text;
            size_t k = 0;
            for (; k < fields.size(); k++) if (name == fields[k].first) break;
            if (k == fields.size()) fail("undefined field: " + name);
            expect_kind(Tok::Colon);
            fields[k].second();
            seen[k] = true;
        }
        for (bool b : seen) if (!b) fail("missing one or more fields");
    }
    // fn_parse_box!
    void parse_box(const std::vector<std::pair<const char*, FieldFn>>& classes) {
        std::string cls = expect_kind(Tok::Identifier).text;
        for (auto& c : classes) if (cls == c.first) { c.second(); return; }
        fail("no such class: " + cls);
    }
    // fn_parse_function!
    void parse_function(const char* name, const std::vector<FieldFn>& params) {
        expect_ident(name);
        expect_kind(Tok::LParen);
        for (size_t i = 0; i < params.size(); i++) {
            if (i) expect_kind(Tok::Comma);
            params[i]();
        }
        expect_kind(Tok::RParen);
    }

    // ---- per-type grammars, src/serialize.rs:606-814 ----
    struct ShapeOut { bool is_plane; V3d a, b; double radius; };
    ShapeOut shape() {
        ShapeOut o{};
        parse_box({{"Sphere", [&] { o.is_plane = false; parse_struct({{"center", [&] { o.a = vec3(); }}, {"radius", [&] { o.radius = f64(); }}}); }},
                   {"Plane", [&] { o.is_plane = true; parse_struct({{"point", [&] { o.a = vec3(); }}, {"normal", [&] { o.b = vec3(); }}}); }}});
        return o;
    }
    MaterialRec material() {
        MaterialRec m{};
        V3d d{}, s{}, am{};
        auto fd = [&] { d = color(); };
        auto fs = [&] { s = color(); };
        auto fe = [&] { m.exponent = f64(); };
        auto fa = [&] { am = color(); };
        parse_box({{"PhongMaterial", [&] { m.kind = MAT_PHONG; m.ior = 1.0; parse_struct({{"diffuse", fd}, {"specular", fs}, {"exponent", fe}, {"ambient", fa}}); }},
                   {"IndirectPhongMaterial", [&] { m.kind = MAT_INDIRECT; m.ior = 1.0; parse_struct({{"diffuse", fd}, {"specular", fs}, {"exponent", fe}, {"ambient", fa}, {"samples", [&] { m.samples = u32(); }}}); }},
                   {"FresnelMaterial", [&] { m.kind = MAT_FRESNEL; parse_struct({{"diffuse", fd}, {"specular", fs}, {"exponent", fe}, {"ambient", fa}, {"ior", [&] { m.ior = f64(); }}}); }},
                   {"TransparentMaterial", [&] { m.kind = MAT_TRANSPARENT; parse_struct({{"specular", fs}, {"exponent", fe}, {"ior", [&] { m.ior = f64(); }}}); }}});
        for (int i = 0; i < 3; i++) { m.diffuse[i] = d.v[i]; m.specular[i] = s.v[i]; m.ambient[i] = am.v[i]; }
        return m;
    }
    void object() {
        ShapeOut sh{};
        MaterialRec m{};
        parse_struct({{"bounds", [&] { sh = shape(); }}, {"material", [&] { m = material(); }}});
        scene_touch(sc);
        sc->materials.push_back(m);
        HostObject o{};
        o.is_plane = sh.is_plane;
        std::memcpy(o.p, sh.a.v, sizeof o.p);
        std::memcpy(o.n, sh.b.v, sizeof o.n);
        o.radius = sh.radius;
        o.material = (uint32_t)sc->materials.size() - 1;
        sc->objects.push_back(o);
        if (o.is_plane) sc->n_planes++; else sc->n_spheres++;
    }
    void light() {
        LightRec l{};
        V3d col{};
        parse_struct({{"model", [&] {
                           parse_box({{"PointLight", [&] { l.kind = LIGHT_POINT; parse_struct({{"location", [&] { V3d v = vec3(); std::memcpy(l.a, v.v, sizeof l.a); }}}); }},
                                      {"DirectionalLight", [&] { l.kind = LIGHT_DIRECTIONAL; parse_struct({{"direction", [&] { V3d v = vec3(); std::memcpy(l.a, v.v, sizeof l.a); }}}); }},
                                      {"AreaLight", [&] {
                                           l.kind = LIGHT_AREA;
                                           parse_struct({{"origin", [&] { V3d v = vec3(); std::memcpy(l.a, v.v, sizeof l.a); }},
                                                         {"side1", [&] { V3d v = vec3(); std::memcpy(l.s1, v.v, sizeof l.s1); }},
                                                         {"side2", [&] { V3d v = vec3(); std::memcpy(l.s2, v.v, sizeof l.s2); }}});
                                       }}});
                       }},
                      {"color", [&] { col = color(); }}});
        std::memcpy(l.color, col.v, sizeof l.color);
        sc->lights.push_back(l);
    }
    void spc_new(CameraRec* cam) {
        V3d p{}, l{}, u{};
        double im = 0;
        parse_function("new", {[&] { p = vec3(); }, [&] { l = vec3(); }, [&] { u = vec3(); }, [&] { im = f64(); }});
        camera_spc_new(cam, p.v, l.v, u.v, im);
    }
    void spc_look_at(CameraRec* cam) {
        V3d f{}, l{}, u{};
        double pov = 0, h = 0;
        parse_function("look_at", {[&] { f = vec3(); }, [&] { l = vec3(); }, [&] { u = vec3(); }, [&] { pov = ang(); }, [&] { h = f64(); }});
        camera_spc_look_at(cam, f.v, l.v, u.v, pov, h);
    }
    void spc(CameraRec* cam) {
        try { spc_new(cam); } catch (const ParseError&) { spc_look_at(cam); }
    }
    void camera() {
        CameraRec cam{};
        parse_box({{"SimplePerspectiveCamera", [&] { spc(&cam); }},
                   {"DepthOfFieldCamera", [&] {
                        double focus = 0, aperture = 0;
                        uint32_t samples = 0;
                        parse_function("new", {[&] { spc(&cam); }, [&] { focus = f64(); }, [&] { aperture = f64(); }, [&] { samples = u32(); }});
                        camera_make_dof(&cam, focus, aperture, samples);
                    }}});
        sc->camera = cam;
        sc->camera_set = true;
    }
    void load_texture(HostTexture* out) {
        std::string path;
        parse_function("load", {[&] { path = str(); }});
        std::string err;
        if (!load_texture_file(path, out, &err)) fail("error loading \"" + path + "\": " + err);
    }
    void background() {
        parse_box({{"SolidColorBackground", [&] {
                        V3d c{};
                        parse_struct({{"color", [&] { c = color(); }}});
                        sc->bg_kind = BG_SOLID;
                        std::memcpy(sc->bg_color, c.v, sizeof sc->bg_color);
                    }},
                   {"SkyboxBackground", [&] {
                        HostTexture f[6];
                        parse_struct({{"px", [&] { load_texture(&f[0]); }}, {"nx", [&] { load_texture(&f[1]); }}, {"py", [&] { load_texture(&f[2]); }},
                                      {"ny", [&] { load_texture(&f[3]); }}, {"pz", [&] { load_texture(&f[4]); }}, {"nz", [&] { load_texture(&f[5]); }}});
                        sc->bg_kind = BG_SKYBOX;
                        scene_touch(sc); for (int i = 0; i < 6; i++) sc->faces[i] = std::move(f[i]);
                    }}});
        sc->background_set = true;
    }
    void options() {
        uint32_t w = 0, h = 0, aa = 0;
        parse_struct({{"width", [&] { w = u32(); }}, {"height", [&] { h = u32(); }}, {"antialias", [&] { aa = u32(); }}});
        sc->width = w; sc->height = h; sc->antialias = aa;
        sc->options_set = true;
    }
    template <typename F>
    void parse_vec(F item) {  // src/serialize.rs:596-604
        expect_kind(Tok::LBracket);
        while (!accept(Tok::RBracket)) item();
    }
    void scene() {
        parse_struct({{"objects", [&] {
                           // a repeated `objects:` field replaces the earlier vector upstream
                           scene_touch(sc); sc->objects.clear(); sc->materials.clear(); sc->n_spheres = sc->n_planes = 0;
                           parse_vec([&] { object(); });
                       }},
                      {"lights", [&] { sc->lights.clear(); parse_vec([&] { light(); }); }},
                      {"camera", [&] { camera(); }},
                      {"background", [&] { background(); }},
                      {"options", [&] { options(); }}});
    }

private:
    bool pull(Token* t) {
        bool ok = lex.next(t);
        row = lex.a.row; col = lex.a.col;  // Processable for Tokenizer, src/serialize.rs:419-424
        return ok;
    }
    Token peeked_;
    bool has_peek_ = false;
};

}  // namespace

int parse_scene_text(const char* text, size_t len, rt_scene* scene, std::string* err) {
    Parser p(text, len, scene);
    bool perr = false;
    ParseError pe{};
    try {
        p.scene();
    } catch (const ParseError& e) {
        perr = true;
        pe = e;
    }
    if (p.lex.failed) pe = p.lex.error;  // lexer error overrides, src/serialize.rs:436-440
    if (perr || p.lex.failed) {
        *err = std::to_string(pe.row) + ":" + std::to_string(pe.col) + ": " + pe.msg;
        return RT_ERR_SYNTAX;
    }
    return RT_OK;
}

}  // namespace rt

extern "C" {

int rt_scene_parse(const char* text, size_t len, rt_scene** out, char* err, size_t errcap) {
    if (!text || !out) { rt::set_error("rt_scene_parse: null argument"); return RT_ERR_INVALID; }
    *out = nullptr;
    rt_scene* s = new rt_scene();
    std::string msg;
    int rc = rt::parse_scene_text(text, len, s, &msg);
    if (rc != RT_OK) {
        rt::set_error(msg);
        if (err && errcap) { std::snprintf(err, errcap, "%s", msg.c_str()); }
        delete s;
        return rc;
    }
    *out = s;
    return RT_OK;
}

// std::str::from_utf8's acceptance test (read_to_string fails on anything else, src/main.rs:22-25): shortest forms only,
// no surrogates, nothing above U+10FFFF
static bool valid_utf8(const std::string& s) {
    const unsigned char* p = (const unsigned char*)s.data();
    const unsigned char* e = p + s.size();
    while (p < e) {
        const unsigned char c = *p;
        if (c < 0x80) { p++; continue; }
        int n;
        unsigned lo = 0x80, hi = 0xBF;
        if (c >= 0xC2 && c <= 0xDF) n = 1;
        else if (c == 0xE0) { n = 2; lo = 0xA0; }
        else if (c >= 0xE1 && c <= 0xEC) n = 2;
        else if (c == 0xED) { n = 2; hi = 0x9F; }
        else if (c >= 0xEE && c <= 0xEF) n = 2;
        else if (c == 0xF0) { n = 3; lo = 0x90; }
        else if (c >= 0xF1 && c <= 0xF3) n = 3;
        else if (c == 0xF4) { n = 3; hi = 0x8F; }
        else return false;
        if (e - p <= n) return false;
        if (p[1] < lo || p[1] > hi) return false;
        for (int k = 2; k <= n; k++)
            if (p[k] < 0x80 || p[k] > 0xBF) return false;
        p += n + 1;
    }
    return true;
}

// File::open + read_to_string + deserialize (src/main.rs:16-29).  RT_ERR_IO messages are what `println!("error: {}", e)`
// shows for the io::Error upstream: "<strerror> (os error <errno>)", or "stream did not contain valid UTF-8".
int rt_scene_parse_file(const char* path, rt_scene** out, char* err, size_t errcap) {
    if (!path || !out) { rt::set_error("rt_scene_parse_file: null argument"); return RT_ERR_INVALID; }
    auto io_fail = [&](const std::string& m) {
        rt::set_error(m);
        if (err && errcap) std::snprintf(err, errcap, "%s", m.c_str());
        return RT_ERR_IO;
    };
    errno = 0;
    FILE* f = std::fopen(path, "rb");
    if (!f) {
        const int en = errno;
        return io_fail(std::string(std::strerror(en)) + " (os error " + std::to_string(en) + ")");
    }
    std::string buf;
    char tmp[65536];
    size_t n;
    while ((n = std::fread(tmp, 1, sizeof tmp, f)) > 0) buf.append(tmp, n);
    if (std::ferror(f)) {
        const int en = errno;
        std::fclose(f);
        return io_fail(std::string(std::strerror(en)) + " (os error " + std::to_string(en) + ")");
    }
    std::fclose(f);
    if (!valid_utf8(buf)) return io_fail("stream did not contain valid UTF-8");
    return rt_scene_parse(buf.data(), buf.size(), out, err, errcap);
}

}  // extern "C"
