// Minimal stand-in for mqt-core's ir/Definitions.hpp (mqt-core v3.1.0 is not
// vendored in the reference and cannot be fetched here). TEST INFRASTRUCTURE:
// exists only so the reference's own src/sc sources compile into oracle/_ref.
// It carries no arithmetic of the routing path.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <utility>
#include <vector>

namespace qc {
using Qubit = std::uint32_t;
using Bit = std::uint64_t;
using fp = double;
enum class Format : std::uint8_t { OpenQASM2, OpenQASM3 };
using Targets = std::vector<Qubit>;

class QuantumRegister {
public:
  QuantumRegister() = default;
  QuantumRegister(Qubit start, std::size_t sz, std::string n)
      : startIndex(start), size(sz), name(std::move(n)) {}
  [[nodiscard]] Qubit getStartIndex() const { return startIndex; }
  [[nodiscard]] std::size_t getSize() const { return size; }
  [[nodiscard]] const std::string& getName() const { return name; }
  [[nodiscard]] std::string toString(Qubit i) const {
    return name + "[" + std::to_string(i - startIndex) + "]";
  }
  Qubit startIndex = 0;
  std::size_t size = 0;
  std::string name;
};
class ClassicalRegister {
public:
  ClassicalRegister() = default;
  ClassicalRegister(Bit start, std::size_t sz, std::string n)
      : startIndex(start), size(sz), name(std::move(n)) {}
  [[nodiscard]] Bit getStartIndex() const { return startIndex; }
  [[nodiscard]] std::size_t getSize() const { return size; }
  [[nodiscard]] const std::string& getName() const { return name; }
  [[nodiscard]] std::string toString(Bit i) const {
    return name + "[" + std::to_string(i - startIndex) + "]";
  }
  Bit startIndex = 0;
  std::size_t size = 0;
  std::string name;
};
using QuantumRegisterMap = std::map<std::string, QuantumRegister>;
using ClassicalRegisterMap = std::map<std::string, ClassicalRegister>;
using QubitIndexToRegisterMap =
    std::map<Qubit, std::pair<QuantumRegister, std::string>>;
using BitIndexToRegisterMap =
    std::map<Bit, std::pair<ClassicalRegister, std::string>>;
} // namespace qc
