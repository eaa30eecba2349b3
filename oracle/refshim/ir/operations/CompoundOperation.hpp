// Stand-in for mqt-core's qc::CompoundOperation.
#pragma once
#include "ir/operations/Operation.hpp"
#include <memory>
#include <vector>

namespace qc {
class CompoundOperation : public Operation {
  std::vector<std::unique_ptr<Operation>> ops;

public:
  CompoundOperation() {
    type = Compound;
    name = "compound";
  }
  CompoundOperation(const CompoundOperation& o) : Operation(o) {
    for (const auto& op : o.ops) {
      ops.emplace_back(op->clone());
    }
  }
  [[nodiscard]] std::unique_ptr<Operation> clone() const override {
    return std::make_unique<CompoundOperation>(*this);
  }
  [[nodiscard]] bool isCompoundOperation() const override { return true; }
  void emplace_back(std::unique_ptr<Operation> op) {
    ops.emplace_back(std::move(op));
  }
  [[nodiscard]] auto begin() { return ops.begin(); }
  [[nodiscard]] auto end() { return ops.end(); }
  [[nodiscard]] auto begin() const { return ops.begin(); }
  [[nodiscard]] auto end() const { return ops.end(); }
  [[nodiscard]] auto cbegin() const { return ops.cbegin(); }
  [[nodiscard]] auto cend() const { return ops.cend(); }
  [[nodiscard]] std::size_t size() const { return ops.size(); }
  void dumpOpenQASM3(std::ostream& of, const QubitIndexToRegisterMap& qreg,
                     const BitIndexToRegisterMap& creg) const {
    for (const auto& op : ops) {
      op->dumpOpenQASM(of, qreg, creg, 0, true);
    }
  }
};
} // namespace qc
