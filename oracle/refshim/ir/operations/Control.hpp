// Stand-in for mqt-core's qc::Control / qc::Controls.
#pragma once
#include "ir/Definitions.hpp"
#include <set>

namespace qc {
struct Control {
  enum class Type : bool { Pos = true, Neg = false };
  Qubit qubit{};
  Type type = Type::Pos;
  // NOLINTNEXTLINE(google-explicit-constructor)
  Control(const Qubit q = {}, const Type t = Type::Pos) : qubit(q), type(t) {}
};
inline bool operator<(const Control& a, const Control& b) {
  return a.qubit < b.qubit || (a.qubit == b.qubit && a.type < b.type);
}
inline bool operator==(const Control& a, const Control& b) {
  return a.qubit == b.qubit && a.type == b.type;
}
using Controls = std::set<Control>;
} // namespace qc
