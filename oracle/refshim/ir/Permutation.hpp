// Stand-in for mqt-core's qc::Permutation (ordered map physical -> logical).
#pragma once
#include "ir/Definitions.hpp"
#include <map>

namespace qc {
class Permutation {
  std::map<Qubit, Qubit> m;

public:
  using iterator = std::map<Qubit, Qubit>::iterator;
  using const_iterator = std::map<Qubit, Qubit>::const_iterator;
  auto begin() { return m.begin(); }
  auto end() { return m.end(); }
  [[nodiscard]] auto begin() const { return m.begin(); }
  [[nodiscard]] auto end() const { return m.end(); }
  [[nodiscard]] auto cbegin() const { return m.cbegin(); }
  [[nodiscard]] auto cend() const { return m.cend(); }
  auto rbegin() { return m.rbegin(); }
  auto rend() { return m.rend(); }
  [[nodiscard]] auto rbegin() const { return m.rbegin(); }
  [[nodiscard]] auto rend() const { return m.rend(); }
  Qubit& at(Qubit k) { return m.at(k); }
  [[nodiscard]] const Qubit& at(Qubit k) const { return m.at(k); }
  Qubit& operator[](Qubit k) { return m[k]; }
  [[nodiscard]] std::size_t count(Qubit k) const { return m.count(k); }
  auto find(Qubit k) { return m.find(k); }
  [[nodiscard]] auto find(Qubit k) const { return m.find(k); }
  [[nodiscard]] std::size_t size() const { return m.size(); }
  [[nodiscard]] bool empty() const { return m.empty(); }
  void clear() { m.clear(); }
  auto erase(Qubit k) { return m.erase(k); }
  auto erase(const_iterator it) { return m.erase(it); }
  template <class... A> auto emplace(A&&... a) {
    return m.emplace(std::forward<A>(a)...);
  }
  auto insert(const std::pair<const Qubit, Qubit>& p) { return m.insert(p); }
};
} // namespace qc
