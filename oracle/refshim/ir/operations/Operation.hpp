// Stand-in for mqt-core's qc::Operation: only the surface the reference's
// src/sc sources call. TEST INFRASTRUCTURE.
#pragma once
#include "ir/Definitions.hpp"
#include "ir/operations/Control.hpp"
#include "ir/operations/OpType.hpp"
#include <memory>
#include <ostream>
#include <string>
#include <vector>

namespace qc {
class Operation {
protected:
  Controls controls;
  Targets targets;
  std::vector<fp> parameter;
  OpType type = None;
  std::string name;

public:
  Operation() = default;
  Operation(const Operation&) = default;
  Operation& operator=(const Operation&) = default;
  virtual ~Operation() = default;
  [[nodiscard]] virtual std::unique_ptr<Operation> clone() const = 0;
  [[nodiscard]] const Targets& getTargets() const { return targets; }
  Targets& getTargets() { return targets; }
  [[nodiscard]] const Controls& getControls() const { return controls; }
  Controls& getControls() { return controls; }
  [[nodiscard]] const std::vector<fp>& getParameter() const {
    return parameter;
  }
  [[nodiscard]] OpType getType() const { return type; }
  [[nodiscard]] const std::string& getName() const { return name; }
  void setTargets(const Targets& t) { targets = t; }
  void setControls(const Controls& c) { controls = c; }
  [[nodiscard]] virtual bool isUnitary() const {
    return type != Measure && type != Reset && type != ClassicControlled;
  }
  [[nodiscard]] virtual bool isStandardOperation() const { return false; }
  [[nodiscard]] virtual bool isCompoundOperation() const { return false; }
  [[nodiscard]] virtual bool isControlled() const { return !controls.empty(); }
  [[nodiscard]] bool actsOn(Qubit q) const {
    for (const auto& t : targets) {
      if (t == q) {
        return true;
      }
    }
    return controls.count(Control{q}) > 0;
  }
  virtual void dumpOpenQASM(std::ostream& of,
                            const QubitIndexToRegisterMap& qreg,
                            const BitIndexToRegisterMap& creg, std::size_t,
                            bool) const {
    (void)creg;
    auto qname = [&](Qubit q) {
      auto it = qreg.find(q);
      return it != qreg.end() ? it->second.second
                              : "q[" + std::to_string(q) + "]";
    };
    if (type == Barrier) {
      of << "barrier";
    } else {
      of << (controls.empty() ? "" : "c") << toString(type);
      if (!parameter.empty()) {
        of << "(";
        for (std::size_t i = 0; i < parameter.size(); ++i) {
          of << (i ? "," : "") << parameter[i];
        }
        of << ")";
      }
    }
    bool first = true;
    for (const auto& c : controls) {
      of << (first ? " " : ", ") << qname(c.qubit);
      first = false;
    }
    for (const auto& t : targets) {
      of << (first ? " " : ", ") << qname(t);
      first = false;
    }
    of << ";\n";
  }
};
} // namespace qc
