// Engine stand-in for the C++ class API shims (include/groth16.hpp, include/ultra_groth.hpp).
//
// The reference's Groth16::Prover<Engine> / UltraGroth::Prover<Engine> are templates over ffiasm's AltBn128::Engine (src/groth16.hpp:10,
// src/prover.cpp:6 -- iden3/ffiasm, absent from the reference tree).  The shims only need the element / point TYPES of that engine,
// with the memory layout the zkey sections are cast to (src/groth16.cpp:33-43): 4x64-bit little-endian limbs, Montgomery form for
// coordinates, raw integers for witness signals; {x, y} affine, all-zero = infinity; Fq2 element {a, b} = a + b*u.
#ifndef UGB_ENGINE_HPP
#define UGB_ENGINE_HPP
#include <cstdint>

namespace UgbAltBn128 {

struct F1Element { uint64_t v[4]; };
struct F2Element { F1Element a, b; };
struct FrElement { uint64_t v[4]; };
struct G1PointAffine { F1Element x, y; };
struct G2PointAffine { F2Element x, y; };
static_assert(sizeof(G1PointAffine) == 64 && sizeof(G2PointAffine) == 128 && sizeof(FrElement) == 32, "zkey point layout");

class Engine {
public:
    typedef UgbAltBn128::F1Element F1Element;
    typedef UgbAltBn128::F2Element F2Element;
    typedef UgbAltBn128::FrElement FrElement;
    typedef UgbAltBn128::G1PointAffine G1PointAffine;
    typedef UgbAltBn128::G2PointAffine G2PointAffine;
    static Engine& instance() { static Engine e; return e; }
};

}  // namespace UgbAltBn128
#endif
