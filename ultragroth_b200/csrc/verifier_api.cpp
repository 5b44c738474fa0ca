// Host verifier behind the reference's verifier C API (src/verifier.h:22-38; include/verifier.h here):
//   groth16_verify / ultra_groth_verify(proof json, inputs json, verification key json) -> 0 valid, 1 invalid, 2 error.
// Follows src/verifier.cpp:15-243 (parsing, protocol/curve checks, error strings), src/groth16.cpp:314-364 and
// src/ultra_groth.cpp:581-648 (verification equations) and the optimal-ate pairing of src/groth16.cpp:366-690 (loop
// count 6u+2, two Frobenius corrections, constants of :302-311).  O(1) work per proof: it stays on the CPU
// (SURVEY.md 8(f) N2); the pairing lives in Fq12 = Fq[w]/(w^12 - 18 w^6 + 82), w^6 = 9 + u, the same model as the
// oracle (oracle/bn254.py), with a plain square-and-multiply final exponentiation by (q^12-1)/r.
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "curve.cuh"
#include "host_ff.hpp"
#include "hostio.hpp"
#include "verifier_consts.hpp"
#include "../../include/verifier.h"

namespace ug {
namespace {

using HG1Affine = Affine<HFq>;
using HG2Affine = Affine<HFq2>;

// ---- a JSON reader for what snarkjs-style proofs / keys contain: objects, arrays, strings, integers ----
struct JValue {
    enum Kind { STR, NUM, ARR, OBJ, OTHER } kind = OTHER;
    std::string s;  // STR: contents, NUM / OTHER: literal text
    std::vector<JValue> arr;
    std::vector<std::pair<std::string, JValue>> obj;
    const JValue& at(const char* key) const {
        if (kind != OBJ) throw std::invalid_argument("json: not an object");
        for (auto& kv : obj) if (kv.first == key) return kv.second;
        throw std::invalid_argument(std::string("json: missing key ") + key);
    }
    const JValue& at(size_t i) const {
        if (kind != ARR || i >= arr.size()) throw std::invalid_argument("json: bad array access");
        return arr[i];
    }
    const std::string& str() const {
        if (kind != STR) throw std::invalid_argument("json: not a string");
        return s;
    }
};

struct JParser {
    const char* p;
    explicit JParser(const char* text) : p(text) {}
    void ws() { while (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r') p++; }
    std::string string_() {
        if (*p != '"') throw std::invalid_argument("json: expected string");
        p++;
        std::string out;
        while (*p && *p != '"') {
            if (*p == '\\') {
                p++;
                if (!*p) break;
                switch (*p) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': throw std::invalid_argument("json: \\u escapes not supported");
                    default: out += *p;
                }
                p++;
            } else {
                out += *p++;
            }
        }
        if (*p != '"') throw std::invalid_argument("json: unterminated string");
        p++;
        return out;
    }
    JValue value(int depth = 0) {
        if (depth > 64) throw std::invalid_argument("json: too deep");
        ws();
        JValue v;
        if (*p == '"') {
            v.kind = JValue::STR;
            v.s = string_();
        } else if (*p == '[') {
            v.kind = JValue::ARR;
            p++;
            ws();
            if (*p == ']') { p++; return v; }
            while (true) {
                v.arr.push_back(value(depth + 1));
                ws();
                if (*p == ',') { p++; continue; }
                if (*p == ']') { p++; break; }
                throw std::invalid_argument("json: expected , or ]");
            }
        } else if (*p == '{') {
            v.kind = JValue::OBJ;
            p++;
            ws();
            if (*p == '}') { p++; return v; }
            while (true) {
                ws();
                std::string k = string_();
                ws();
                if (*p != ':') throw std::invalid_argument("json: expected :");
                p++;
                v.obj.emplace_back(k, value(depth + 1));
                ws();
                if (*p == ',') { p++; continue; }
                if (*p == '}') { p++; break; }
                throw std::invalid_argument("json: expected , or }");
            }
        } else {
            const char* q = p;
            while (*p && *p != ',' && *p != ']' && *p != '}' && *p != ' ' && *p != '\n' && *p != '\t' && *p != '\r') p++;
            if (p == q) throw std::invalid_argument("json: unexpected character");
            v.s.assign(q, p);
            v.kind = ((*q >= '0' && *q <= '9') || *q == '-') ? JValue::NUM : JValue::OTHER;
        }
        return v;
    }
    JValue parse() {
        JValue v = value();
        ws();
        if (*p) throw std::invalid_argument("json: trailing characters");
        return v;
    }
};

// decimal string -> 256-bit integer reduced mod the field prime (RawFq::fromString reduces mod q, build/fq.cpp), Montgomery
template <class HF>
HF field_from_dec(const std::string& s) {
    if (s.empty()) throw std::invalid_argument("empty number");
    // Horner in the field itself: acc = acc*10 + digit (all in Montgomery form), which reduces as it goes
    HF ten = HF::zero(), acc = HF::zero();
    ten.v[0] = 10;
    ten = to_mont(ten);
    for (char c : s) {
        if (c < '0' || c > '9') throw std::invalid_argument("not a decimal number");
        HF d = HF::zero();
        d.v[0] = (uint64_t)(c - '0');
        acc = fadd(fmul(acc, ten), to_mont(d));
    }
    return acc;
}

HG1Affine g1_from_json(const JValue& v) {
    HG1Affine p;
    p.x = field_from_dec<HFq>(v.at((size_t)0).str());
    p.y = field_from_dec<HFq>(v.at((size_t)1).str());
    return p;
}
HG2Affine g2_from_json(const JValue& v) {
    HG2Affine p;
    p.x.a = field_from_dec<HFq>(v.at((size_t)0).at((size_t)0).str());
    p.x.b = field_from_dec<HFq>(v.at((size_t)0).at((size_t)1).str());
    p.y.a = field_from_dec<HFq>(v.at((size_t)1).at((size_t)0).str());
    p.y.b = field_from_dec<HFq>(v.at((size_t)1).at((size_t)1).str());
    return p;
}

// ---- Fq12 = Fq[w]/(w^12 - 18 w^6 + 82) ----------------------------------------------------------------------
struct F12 {
    HFq c[12];
    static F12 one() { F12 r; for (auto& x : r.c) x = HFq::zero(); r.c[0] = HFq::one(); return r; }
    bool operator==(const F12& o) const { for (int i = 0; i < 12; i++) if (!(c[i] == o.c[i])) return false; return true; }
};
HFq small_fq(uint64_t k) { HFq r = HFq::zero(); r.v[0] = k; return to_mont(r); }

F12 f12_mul(const F12& a, const F12& b) {
    static const HFq c18 = small_fq(18), c82 = small_fq(82);
    HFq t[23];
    for (auto& x : t) x = HFq::zero();
    for (int i = 0; i < 12; i++) {
        if (a.c[i].is_zero()) continue;
        for (int j = 0; j < 12; j++) {
            if (b.c[j].is_zero()) continue;
            t[i + j] = fadd(t[i + j], fmul(a.c[i], b.c[j]));
        }
    }
    for (int k = 22; k >= 12; k--) {  // w^12 = 18 w^6 - 82
        if (t[k].is_zero()) continue;
        t[k - 6] = fadd(t[k - 6], fmul(t[k], c18));
        t[k - 12] = fsub(t[k - 12], fmul(t[k], c82));
    }
    F12 r;
    for (int i = 0; i < 12; i++) r.c[i] = t[i];
    return r;
}

F12 f12_pow(const F12& a, const uint64_t* e, int bits) {
    F12 r = F12::one();
    for (int i = bits - 1; i >= 0; i--) {
        r = f12_mul(r, r);
        if ((e[i >> 6] >> (i & 63)) & 1) r = f12_mul(r, a);
    }
    return r;
}

// a + b*u  ->  (a - 9b) + b*w^6
void embed_f2(const HFq2& x, HFq& c0, HFq& c6) {
    static const HFq nine = small_fq(9);
    c0 = fsub(x.a, fmul(nine, x.b));
    c6 = x.b;
}
HFq2 f2_conj(const HFq2& x) { HFq2 r; r.a = x.a; r.b = fneg(x.b); return r; }
HFq2 f2_from_dec(const char* a, const char* b) { HFq2 r; r.a = field_from_dec<HFq>(a); r.b = field_from_dec<HFq>(b); return r; }

// value at P of the line through the twisted point (x1 w^2, y1 w^3) with slope m*w:  m*w*(xp - x1 w^2) - (yp - y1 w^3)
F12 line_at(const HFq2& m, const HFq2& x1, const HFq2& y1, const HG1Affine& P) {
    F12 o;
    for (auto& x : o.c) x = HFq::zero();
    HFq me0, me6, mx0, mx6, ye0, ye6;
    embed_f2(m, me0, me6);
    embed_f2(fmul(m, x1), mx0, mx6);
    embed_f2(y1, ye0, ye6);
    o.c[0] = fneg(P.y);
    o.c[1] = fmul(me0, P.x);
    o.c[7] = fmul(me6, P.x);
    o.c[3] = fsub(ye0, mx0);
    o.c[9] = fsub(ye6, mx6);
    return o;
}
// xp - x1 w^2 (the two points are opposite)
F12 vertical_at(const HFq2& x1, const HG1Affine& P) {
    F12 o;
    for (auto& x : o.c) x = HFq::zero();
    HFq xe0, xe6;
    embed_f2(x1, xe0, xe6);
    o.c[0] = P.x;
    o.c[2] = fneg(xe0);
    o.c[8] = fneg(xe6);
    return o;
}

struct G2Step { HG2Affine T; HFq2 m; bool inf; };
G2Step g2_dbl(const HG2Affine& T) {
    G2Step s;
    HFq2 xx = fsqr(T.x);
    s.m = fmul(fadd(fdbl(xx), xx), finv(fdbl(T.y)));
    s.T.x = fsub(fsqr(s.m), fdbl(T.x));
    s.T.y = fsub(fmul(s.m, fsub(T.x, s.T.x)), T.y);
    s.inf = false;
    return s;
}
G2Step g2_add(const HG2Affine& T, const HG2Affine& S) {
    if (T.x == S.x) {
        if (T.y == S.y) return g2_dbl(T);
        G2Step s;
        s.inf = true;
        return s;
    }
    G2Step s;
    s.m = fmul(fsub(S.y, T.y), finv(fsub(S.x, T.x)));
    s.T.x = fsub(fsub(fsqr(s.m), T.x), S.x);
    s.T.y = fsub(fmul(s.m, fsub(T.x, s.T.x)), T.y);
    s.inf = false;
    return s;
}

// Miller loop of the optimal ate pairing (src/groth16.cpp:522-609): affine G2 point Q, affine G1 point P, both finite
F12 miller_loop(const HG2Affine& Q, const HG1Affine& P) {
    static const HFq2 xi_p1_3 = f2_from_dec("21575463638280843010398324269430826099269044274347216827212613867836435027261",
                                            "10307601595873709700152284273816112264069230130616436755625194854815875713954");
    static const HFq2 xi_p1_2 = f2_from_dec("2821565182194536844548159561693502659359617185244120367078079554186484126554",
                                            "3505843767911556378687030309984248845540243509899259641013678093033130930403");
    static const HFq xi_p2_3 = field_from_dec<HFq>("21888242871839275220042445260109153167277707414472061641714758635765020556616");
    HG2Affine T = Q;
    F12 f = F12::one();
    for (int i = ATE_BITS - 2; i >= 0; i--) {
        G2Step s = g2_dbl(T);
        f = f12_mul(f12_mul(f, f), line_at(s.m, T.x, T.y, P));
        T = s.T;
        if ((ATE_LOOP[i >> 6] >> (i & 63)) & 1) {
            s = g2_add(T, Q);
            if (s.inf) throw std::invalid_argument("pairing: degenerate G2 point");
            f = f12_mul(f, line_at(s.m, T.x, T.y, P));
            T = s.T;
        }
    }
    HG2Affine Q1, nQ2;
    Q1.x = fmul(f2_conj(Q.x), xi_p1_3);
    Q1.y = fmul(f2_conj(Q.y), xi_p1_2);
    nQ2.x.a = fmul(Q.x.a, xi_p2_3);
    nQ2.x.b = fmul(Q.x.b, xi_p2_3);
    nQ2.y = Q.y;
    G2Step s = g2_add(T, Q1);
    if (s.inf) throw std::invalid_argument("pairing: degenerate G2 point");
    f = f12_mul(f, line_at(s.m, T.x, T.y, P));
    T = s.T;
    s = g2_add(T, nQ2);
    if (s.inf) f = f12_mul(f, vertical_at(T.x, P));
    else f = f12_mul(f, line_at(s.m, T.x, T.y, P));
    return f;
}

// prod e(P_i, Q_i) == 1 (src/groth16.cpp:672-690); pairs with a point at infinity contribute 1
bool pairing_check(const std::vector<HG1Affine>& g1, const std::vector<HG2Affine>& g2) {
    F12 acc = F12::one();
    for (size_t i = 0; i < g1.size(); i++) {
        if (g1[i].is_inf() || g2[i].is_inf()) continue;
        acc = f12_mul(acc, miller_loop(g2[i], g1[i]));
    }
    return f12_pow(acc, FINAL_EXP, FINAL_EXP_BITS) == F12::one();
}

bool g1_on_curve(const HG1Affine& p) {
    if (p.is_inf()) return true;
    static const HFq three = small_fq(3);
    return fsqr(p.y) == fadd(fmul(fsqr(p.x), p.x), three);
}
bool g2_on_curve(const HG2Affine& p) {
    if (p.is_inf()) return true;
    // twist: y^2 = x^3 + 3/(9+u)
    static const HFq2 b2 = [] {
        HFq2 xi; xi.a = small_fq(9); xi.b = HFq::one();
        HFq2 three; three.a = small_fq(3); three.b = HFq::zero();
        return fmul(three, finv(xi));
    }();
    return fsqr(p.y) == fadd(fmul(fsqr(p.x), p.x), b2);
}

// ---- verification equations -----------------------------------------------------------------------------------
XYZZ<HFq> g1_mul(const HG1Affine& p, const HFr& k_raw) {  // E.g1.mulByScalar on the raw (non-Montgomery) scalar
    XYZZ<HFq> base = XYZZ<HFq>::from_affine(p), acc = XYZZ<HFq>::inf();
    for (int i = 255; i >= 0; i--) {
        acc = pt_dbl(acc);
        if ((k_raw.v[i >> 6] >> (i & 63)) & 1) pt_add(acc, base);
    }
    return acc;
}
HG1Affine g1_neg(const HG1Affine& p) { HG1Affine r = p; r.y = fneg(p.y); return r; }

// src/ultra_groth.cpp:32-58: keccak256 of the 32-byte big-endian x || y, digest read big-endian, one MMul by R^2
HFr derive_challenge(const HG1Affine& p) {
    uint8_t buf[64], digest[32];
    HFq x = from_mont(p.x), y = from_mont(p.y);
    for (int i = 0; i < 32; i++) {
        buf[i] = (uint8_t)(x.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
        buf[32 + i] = (uint8_t)(y.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
    }
    keccak256(buf, 64, digest);
    HFr raw = HFr::zero();
    for (int i = 0; i < 32; i++) raw.v[(31 - i) >> 3] |= (uint64_t)digest[i] << (8 * ((31 - i) & 7));
    return to_mont(raw);
}

std::vector<HFr> parse_inputs(const char* inputs_str) {  // src/verifier.cpp:62-91
    std::vector<HFr> inputs;
    try {
        if (!inputs_str) throw std::invalid_argument("null");
        JValue j = JParser(inputs_str).parse();
        if (j.kind != JValue::ARR || j.arr.empty()) throw std::invalid_argument("invalid inputs data");
        for (auto& e : j.arr) inputs.push_back(field_from_dec<HFr>(e.str()));
    } catch (...) {
        throw std::invalid_argument("invalid inputs data");
    }
    return inputs;
}

struct KeyCommon {
    HG1Affine alpha1;
    HG2Affine beta2, gamma2;
    std::vector<HG1Affine> IC;
};
void parse_key_common(const JValue& k, const char* protocol, KeyCommon& out) {
    if (k.at("protocol").str() != protocol || k.at("curve").str() != "bn128") throw std::invalid_argument("invalid verification key data");
    const JValue& np = k.at("nPublic");
    if (np.kind != JValue::NUM) throw std::invalid_argument("invalid verification key data");
    out.alpha1 = g1_from_json(k.at("vk_alpha_1"));
    out.beta2 = g2_from_json(k.at("vk_beta_2"));
    out.gamma2 = g2_from_json(k.at("vk_gamma_2"));
    const JValue& ic = k.at("IC");
    if (ic.kind != JValue::ARR) throw std::invalid_argument("invalid verification key data");
    for (auto& e : ic.arr) out.IC.push_back(g1_from_json(e));
    if (out.IC.empty()) throw std::invalid_argument("invalid verification key data");
}

bool verify_groth16(const char* proof_str, const char* inputs_str, const char* key_str) {
    HG1Affine A, C;
    HG2Affine B;
    try {  // src/verifier.cpp:15-36
        if (!proof_str) throw std::invalid_argument("null");
        JValue j = JParser(proof_str).parse();
        if (j.at("protocol").str() != "groth16") throw std::invalid_argument("invalid proof data");
        A = g1_from_json(j.at("pi_a"));
        B = g2_from_json(j.at("pi_b"));
        C = g1_from_json(j.at("pi_c"));
    } catch (...) {
        throw std::invalid_argument("invalid proof data");
    }
    std::vector<HFr> inputs = parse_inputs(inputs_str);
    KeyCommon key;
    HG2Affine delta2;
    try {  // src/verifier.cpp:93-120
        if (!key_str) throw std::invalid_argument("null");
        JValue k = JParser(key_str).parse();
        parse_key_common(k, "groth16", key);
        delta2 = g2_from_json(k.at("vk_delta_2"));
    } catch (...) {
        throw std::invalid_argument("invalid verification key data");
    }
    // src/groth16.cpp:314-364
    if (inputs.size() + 1 != key.IC.size()) throw std::invalid_argument("len(inputs)+1 != len(vk.IC)");
    if (!g1_on_curve(A) || !g1_on_curve(C) || !g2_on_curve(B)) return false;
    XYZZ<HFq> vkx = XYZZ<HFq>::inf();
    for (size_t i = 0; i < inputs.size(); i++) pt_add(vkx, g1_mul(key.IC[i + 1], from_mont(inputs[i])));
    pt_add(vkx, XYZZ<HFq>::from_affine(key.IC[0]));
    return pairing_check({A, g1_neg(key.alpha1), g1_neg(pt_to_affine(vkx)), g1_neg(C)}, {B, key.beta2, key.gamma2, delta2});
}

bool verify_ultragroth(const char* proof_str, const char* inputs_str, const char* key_str) {
    HG1Affine A, Fc, Rc;
    HG2Affine B;
    try {  // src/verifier.cpp:38-60, src/ultra_groth.cpp:533-540
        if (!proof_str) throw std::invalid_argument("null");
        JValue j = JParser(proof_str).parse();
        if (j.at("protocol").str() != "ultragroth") throw std::invalid_argument("invalid proof data");
        A = g1_from_json(j.at("pi_a"));
        B = g2_from_json(j.at("pi_b"));
        Fc = g1_from_json(j.at("pi_f"));
        Rc = g1_from_json(j.at("pi_r"));
    } catch (...) {
        throw std::invalid_argument("invalid proof data");
    }
    std::vector<HFr> inputs = parse_inputs(inputs_str);
    KeyCommon key;
    HG2Affine final_delta2, round_delta2;
    HG1Affine ic_rand;
    try {  // src/verifier.cpp:122-147, src/ultra_groth.cpp:542-563
        if (!key_str) throw std::invalid_argument("null");
        JValue k = JParser(key_str).parse();
        parse_key_common(k, "ultragroth", key);
        final_delta2 = g2_from_json(k.at("vk_delta_c2_2"));
        round_delta2 = g2_from_json(k.at("vk_delta_c1_2"));
        ic_rand = g1_from_json(k.at("IC_rand"));
    } catch (...) {
        throw std::invalid_argument("invalid verification key data");
    }
    // src/ultra_groth.cpp:581-648
    if (inputs.size() + 1 != key.IC.size()) throw std::invalid_argument("len(inputs) != len(vk.IC)");
    if (!g1_on_curve(A) || !g1_on_curve(Fc) || !g1_on_curve(Rc) || !g2_on_curve(B)) return false;
    XYZZ<HFq> vkx = XYZZ<HFq>::inf();
    for (size_t i = 0; i < inputs.size(); i++) pt_add(vkx, g1_mul(key.IC[i + 1], from_mont(inputs[i])));
    XYZZ<HFq> pr = g1_mul(ic_rand, from_mont(derive_challenge(Rc)));
    pt_add(vkx, XYZZ<HFq>::from_affine(key.IC[0]));
    pt_add(vkx, pr);
    return pairing_check({A, g1_neg(key.alpha1), g1_neg(pt_to_affine(vkx)), g1_neg(Fc), g1_neg(Rc)},
                         {B, key.beta2, key.gamma2, final_delta2, round_delta2});
}

template <class Fn>
int guarded(Fn fn, char* error_msg, unsigned long error_msg_maxsize) {  // src/verifier.cpp:157-196
    try {
        return fn() ? VERIFIER_VALID_PROOF : VERIFIER_INVALID_PROOF;
    } catch (std::exception& e) {
        if (error_msg) strncpy(error_msg, e.what(), error_msg_maxsize);
        return VERIFIER_ERROR;
    } catch (...) {
        if (error_msg) strncpy(error_msg, "unknown error", error_msg_maxsize);
        return VERIFIER_ERROR;
    }
}

}  // namespace
}  // namespace ug

extern "C" {
int groth16_verify(const char* proof, const char* inputs, const char* verification_key, char* error_msg, unsigned long error_msg_maxsize) {
    return ug::guarded([&] { return ug::verify_groth16(proof, inputs, verification_key); }, error_msg, error_msg_maxsize);
}
int ultra_groth_verify(const char* proof, const char* inputs, const char* verification_key, char* error_msg, unsigned long error_msg_maxsize) {
    return ug::guarded([&] { return ug::verify_ultragroth(proof, inputs, verification_key); }, error_msg, error_msg_maxsize);
}
}
