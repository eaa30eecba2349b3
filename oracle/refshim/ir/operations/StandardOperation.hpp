// Stand-in for mqt-core's qc::StandardOperation.
#pragma once
#include "ir/operations/Operation.hpp"

namespace qc {
class StandardOperation : public Operation {
public:
  StandardOperation() = default;
  StandardOperation(Qubit target, OpType g, std::vector<fp> params = {}) {
    type = g;
    targets = {target};
    parameter = std::move(params);
    name = toString(g);
  }
  StandardOperation(const Targets& targ, OpType g,
                    std::vector<fp> params = {}) {
    type = g;
    targets = targ;
    parameter = std::move(params);
    name = toString(g);
  }
  StandardOperation(Control control, Qubit target, OpType g,
                    const std::vector<fp>& params = {})
      : StandardOperation(target, g, params) {
    controls.insert(control);
  }
  StandardOperation(const Controls& c, Qubit target, OpType g,
                    const std::vector<fp>& params = {})
      : StandardOperation(target, g, params) {
    controls = c;
  }
  StandardOperation(Qubit target0, Qubit target1, OpType g,
                    const std::vector<fp>& params = {})
      : StandardOperation(Targets{target0, target1}, g, params) {}
  [[nodiscard]] std::unique_ptr<Operation> clone() const override {
    return std::make_unique<StandardOperation>(*this);
  }
  [[nodiscard]] bool isStandardOperation() const override { return true; }
};

// measurement stand-in (non-unitary); only used to exercise the
// "strip final measurements" / "non-unitary" branches
class NonUnitaryOperation : public Operation {
public:
  std::vector<Bit> classics;
  NonUnitaryOperation(const Targets& q, std::vector<Bit> c)
      : classics(std::move(c)) {
    type = Measure;
    targets = q;
    name = "measure";
  }
  [[nodiscard]] std::unique_ptr<Operation> clone() const override {
    return std::make_unique<NonUnitaryOperation>(*this);
  }
  [[nodiscard]] bool isUnitary() const override { return false; }
};
} // namespace qc
