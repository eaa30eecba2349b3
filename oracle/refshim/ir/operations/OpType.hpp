// Stand-in for mqt-core's ir/operations/OpType.hpp: the operation tags the
// sc sources reference, with string conversion.
#pragma once
#include <cstdint>
#include <ostream>
#include <stdexcept>
#include <string>

namespace qc {
enum OpType : std::uint8_t {
  None, GPhase, I, Barrier, H, X, Y, Z, S, Sdg, T, Tdg, V, Vdg, U, U2, P, SX,
  SXdg, RX, RY, RZ, SWAP, iSWAP, Peres, Peresdg, DCX, ECR, RXX, RYY, RZZ, RZX,
  XXminusYY, XXplusYY, Compound, Measure, Reset, ClassicControlled, OpTypeEnd
};

inline std::string toString(const OpType t) {
  switch (t) {
  case None: return "none";
  case GPhase: return "gphase";
  case I: return "id";
  case Barrier: return "barrier";
  case H: return "h";
  case X: return "x";
  case Y: return "y";
  case Z: return "z";
  case S: return "s";
  case Sdg: return "sdg";
  case T: return "t";
  case Tdg: return "tdg";
  case V: return "v";
  case Vdg: return "vdg";
  case U: return "u3";
  case U2: return "u2";
  case P: return "p";
  case SX: return "sx";
  case SXdg: return "sxdg";
  case RX: return "rx";
  case RY: return "ry";
  case RZ: return "rz";
  case SWAP: return "swap";
  case iSWAP: return "iswap";
  case Peres: return "peres";
  case Peresdg: return "peresdg";
  case DCX: return "dcx";
  case ECR: return "ecr";
  case RXX: return "rxx";
  case RYY: return "ryy";
  case RZZ: return "rzz";
  case RZX: return "rzx";
  case XXminusYY: return "xx_minus_yy";
  case XXplusYY: return "xx_plus_yy";
  case Compound: return "compound";
  case Measure: return "measure";
  case Reset: return "reset";
  case ClassicControlled: return "classic_controlled";
  default: return "unknown";
  }
}

inline OpType opTypeFromString(const std::string& s) {
  if (s == "cx" || s == "cnot") {
    return X;
  }
  if (s == "u1" || s == "phase") {
    return P;
  }
  if (s == "u" || s == "u3") {
    return U;
  }
  if (s == "i" || s == "id") {
    return I;
  }
  for (std::uint8_t t = 0; t < OpTypeEnd; ++t) {
    if (toString(static_cast<OpType>(t)) == s) {
      return static_cast<OpType>(t);
    }
  }
  throw std::invalid_argument("Unknown operation type: " + s);
}

inline std::ostream& operator<<(std::ostream& os, const OpType t) {
  return os << toString(t);
}
} // namespace qc
