// qasm_reader.hpp - OpenQASM 2 reader for the gate set of the reference's
// examples/*.qasm (qreg / creg, cx, one-qubit gates with parameter
// expressions, measure, barrier) and the view Mapper's constructor takes of
// the circuit (src/sc/Mapper.cpp:33-53: idle qubits stripped, operands read
// through initialLayout).  The reference delegates reading to mqt-core's
// qasm3::Importer::importf (test/sc/heuristic/test_heuristic.cpp:253), which
// is not part of it; this is the host C++ side of SURVEY section 8 row f2 for
// inputs of 10^5..10^6 gates.  Same results as qmap_b200/circuit.py
// (tests/test_host.py compares them on every example circuit).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace qmg {

// gate kinds; the order is the ABI (qmg_qasm_gate_name)
inline const char* const* qasmGateNames(size_t* count) {
  static const char* const kNames[] = {
      "cx", "h",  "x",  "y",  "z",  "s",  "sdg", "t",  "tdg", "id", "sx",
      "sxdg", "rx", "ry", "rz", "u1", "u2", "u3", "u", "p", "swap"};
  if (count != nullptr) {
    *count = sizeof(kNames) / sizeof(kNames[0]);
  }
  return kNames;
}
constexpr uint8_t kQasmCx = 0;
constexpr uint8_t kQasmSwap = 20;

struct QasmCircuit {
  uint32_t nQubits = 0, nClbits = 0;
  std::vector<int32_t> control, target;  // control = -1 for one-qubit gates
  std::vector<uint8_t> kind;
  std::vector<double> params;  // 3 per gate, unused ones 0
  std::vector<uint8_t> nParams;
  std::vector<uint32_t> measQubit, measClbit;

  // ---- parameter expressions: numbers, pi, + - * /, parentheses, unary sign;
  // evaluated left to right like qmap_b200/circuit.py::_eval_expr
  struct Expr {
    const char* p;
    const char* end;
    void skip() {
      while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) {
        ++p;
      }
    }
    [[noreturn]] void bad() const {
      throw std::invalid_argument("cannot parse parameter expression");
    }
    double atom() {
      skip();
      if (p >= end) {
        bad();
      }
      if (*p == '(') {
        ++p;
        const double v = expr();
        skip();
        if (p >= end || *p != ')') {
          bad();
        }
        ++p;
        return v;
      }
      if (*p == '-') {
        ++p;
        return -atom();
      }
      if (*p == '+') {
        ++p;
        return atom();
      }
      if (end - p >= 2 && p[0] == 'p' && p[1] == 'i') {
        p += 2;
        return 3.141592653589793;  // math.pi
      }
      if ((*p >= '0' && *p <= '9') || *p == '.') {
        // [0-9]*\.?[0-9]+([eE][-+]?[0-9]+)?
        const char* q = p;
        while (q < end && *q >= '0' && *q <= '9') {
          ++q;
        }
        if (q < end && *q == '.') {
          ++q;
        }
        while (q < end && *q >= '0' && *q <= '9') {
          ++q;
        }
        if (q < end && (*q == 'e' || *q == 'E')) {
          const char* r = q + 1;
          if (r < end && (*r == '-' || *r == '+')) {
            ++r;
          }
          if (r < end && *r >= '0' && *r <= '9') {
            while (r < end && *r >= '0' && *r <= '9') {
              ++r;
            }
            q = r;
          }
        }
        const std::string tok(p, q);
        char* stop = nullptr;
        const double v = std::strtod(tok.c_str(), &stop);
        if (stop == tok.c_str() || *stop != '\0') {
          bad();
        }
        p = q;
        return v;
      }
      bad();
    }
    double term() {
      double v = atom();
      for (;;) {
        skip();
        if (p < end && (*p == '*' || *p == '/')) {
          const char o = *p++;
          const double r = atom();
          v = o == '*' ? v * r : v / r;
        } else {
          return v;
        }
      }
    }
    double expr() {
      double v = term();
      for (;;) {
        skip();
        if (p < end && (*p == '+' || *p == '-')) {
          const char o = *p++;
          const double r = term();
          v = o == '+' ? v + r : v - r;
        } else {
          return v;
        }
      }
    }
  };
  static double evalExpr(const char* b, const char* e) {
    Expr x{b, e};
    const double v = x.expr();
    x.skip();
    if (x.p != e) {
      x.bad();
    }
    return v;
  }

  // ---- statements
  struct Reg {
    uint32_t first, size;
  };
  static bool isSpace(char c) {
    return c == ' ' || c == '\t' || c == '\n' || c == '\r';
  }
  static bool isWord(char c) {
    return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') ||
           (c >= '0' && c <= '9') || c == '_';
  }
  static void trim(const char*& b, const char*& e) {
    while (b < e && isSpace(*b)) {
      ++b;
    }
    while (e > b && isSpace(e[-1])) {
      --e;
    }
  }
  [[noreturn]] static void badStatement(const char* b, const char* e) {
    throw std::invalid_argument("cannot parse statement: '" +
                                std::string(b, std::min<size_t>(e - b, 80)) + "'");
  }
  // name [ index ]  -> (register, index) ; index = -1 without brackets
  static std::pair<std::string, long> regArg(const char* b, const char* e) {
    trim(b, e);
    const char* p = b;
    while (p < e && isWord(*p)) {
      ++p;
    }
    if (p == b) {
      throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
    }
    const std::string name(b, p);
    while (p < e && isSpace(*p)) {
      ++p;
    }
    if (p == e) {
      return {name, -1};
    }
    if (*p != '[') {
      throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
    }
    ++p;
    while (p < e && isSpace(*p)) {
      ++p;
    }
    const char* d = p;
    long idx = 0;
    while (p < e && *p >= '0' && *p <= '9') {
      if (idx > 100000000L) {  // (far beyond any register; avoids overflow)
        throw std::invalid_argument("index too large: '" + std::string(b, e) + "'");
      }
      idx = idx * 10 + (*p - '0');
      ++p;
    }
    if (p == d) {
      throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
    }
    while (p < e && isSpace(*p)) {
      ++p;
    }
    if (p >= e || *p != ']') {
      throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
    }
    ++p;
    while (p < e && isSpace(*p)) {
      ++p;
    }
    if (p != e) {
      throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
    }
    return {name, idx};
  }

  void parse(const char* text, size_t len) {
    // comments out
    std::string src;
    src.reserve(len);
    for (size_t i = 0; i < len; ++i) {
      if (text[i] == '/' && i + 1 < len && text[i + 1] == '/') {
        while (i < len && text[i] != '\n') {
          ++i;
        }
        if (i < len) {
          src.push_back('\n');
        }
        continue;
      }
      src.push_back(text[i]);
    }
    std::map<std::string, Reg> qregs, cregs;
    std::map<std::string, uint8_t> kinds;
    {
      size_t cnt = 0;
      const char* const* names = qasmGateNames(&cnt);
      for (size_t k = 0; k < cnt; ++k) {
        kinds[names[k]] = static_cast<uint8_t>(k);
      }
      kinds["CX"] = kQasmCx;
    }
    auto qubitOf = [&](const char* b, const char* e) -> int32_t {
      const auto [name, idx] = regArg(b, e);
      const auto it = qregs.find(name);
      if (it == qregs.end() || idx < 0 || idx >= static_cast<long>(it->second.size)) {
        throw std::invalid_argument("bad qubit argument: '" + std::string(b, e) + "'");
      }
      return static_cast<int32_t>(it->second.first + idx);
    };
    const char* p = src.data();
    const char* const end = p + src.size();
    while (p < end) {
      const char* semi = static_cast<const char*>(std::memchr(p, ';', end - p));
      const char* b = p;
      const char* e = semi != nullptr ? semi : end;
      p = semi != nullptr ? semi + 1 : end;
      trim(b, e);
      if (b == e) {
        continue;
      }
      // first word
      const char* w = b;
      while (w < e && isWord(*w)) {
        ++w;
      }
      const std::string head(b, w);
      if (head == "OPENQASM" || head == "include" || head == "barrier") {
        continue;
      }
      if (head == "qreg" || head == "creg") {
        const auto [name, size] = regArg(w, e);
        if (size < 0) {
          badStatement(b, e);
        }
        if (static_cast<uint64_t>(head == "qreg" ? nQubits : nClbits) + static_cast<uint64_t>(size) >
            (1ULL << 24)) {
          throw std::invalid_argument("register too large: '" + std::string(b, e) + "'");
        }
        if (head == "qreg") {
          qregs[name] = Reg{nQubits, static_cast<uint32_t>(size)};
          nQubits += static_cast<uint32_t>(size);
        } else {
          cregs[name] = Reg{nClbits, static_cast<uint32_t>(size)};
          nClbits += static_cast<uint32_t>(size);
        }
        continue;
      }
      if (head == "measure") {
        const char* arrow = nullptr;
        for (const char* q = w; q + 1 < e; ++q) {
          if (q[0] == '-' && q[1] == '>') {
            arrow = q;
            break;
          }
        }
        if (arrow == nullptr) {
          throw std::invalid_argument("bad measure: '" + std::string(b, e) + "'");
        }
        const auto [cname, cidx] = regArg(arrow + 2, e);
        const auto cit = cregs.find(cname);
        if (cit == cregs.end()) {
          throw std::invalid_argument("bad measure: '" + std::string(b, e) + "'");
        }
        if (cidx < 0) {
          const auto [qname, qidx] = regArg(w, arrow);
          const auto qit = qregs.find(qname);
          if (qit == qregs.end() || qidx >= 0) {
            throw std::invalid_argument("bad measure: '" + std::string(b, e) + "'");
          }
          if (qit->second.size > cit->second.size) {
            throw std::invalid_argument("bad measure: '" + std::string(b, e) + "'");
          }
          for (uint32_t k = 0; k < qit->second.size; ++k) {
            measQubit.push_back(qit->second.first + k);
            measClbit.push_back(cit->second.first + k);
          }
        } else {
          if (cidx >= static_cast<long>(cit->second.size)) {
            throw std::invalid_argument("bad measure: '" + std::string(b, e) + "'");
          }
          measQubit.push_back(static_cast<uint32_t>(qubitOf(w, arrow)));
          measClbit.push_back(cit->second.first + static_cast<uint32_t>(cidx));
        }
        continue;
      }
      // gate [ ( params ) ] args
      if (head.empty()) {
        badStatement(b, e);
      }
      const char* q = w;
      while (q < e && isSpace(*q)) {
        ++q;
      }
      double pv[3] = {0., 0., 0.};
      uint8_t np = 0;
      if (q < e && *q == '(') {
        // matching parenthesis
        int depth = 0;
        const char* r = q;
        for (; r < e; ++r) {
          depth += *r == '(';
          depth -= *r == ')';
          if (depth == 0) {
            break;
          }
        }
        if (r >= e) {
          badStatement(b, e);
        }
        // split at top-level commas
        const char* s = q + 1;
        int d2 = 0;
        for (const char* t = q + 1; t <= r; ++t) {
          if (t == r || (*t == ',' && d2 == 0)) {
            const char* sb = s;
            const char* se = t;
            trim(sb, se);
            if (sb != se) {
              if (np >= 3) {
                badStatement(b, e);
              }
              pv[np++] = evalExpr(sb, se);
            }
            s = t + 1;
            continue;
          }
          d2 += *t == '(';
          d2 -= *t == ')';
        }
        q = r + 1;
        if (q >= e || !isSpace(*q)) {
          badStatement(b, e);
        }
      } else if (q == w) {
        badStatement(b, e);  // (no blank between the gate and its arguments)
      }
      // arguments
      int32_t qs[2] = {-1, -1};
      int nArgs = 0;
      const char* s = q;
      for (const char* t = q; t <= e; ++t) {
        if (t == e || *t == ',') {
          if (nArgs < 2) {
            qs[nArgs] = qubitOf(s, t);
          }
          ++nArgs;
          s = t + 1;
        }
      }
      const auto kit = kinds.find(head);
      const bool known = kit != kinds.end();
      if (known && kit->second == kQasmCx) {
        if (nArgs != 2) {
          throw std::invalid_argument("bad cx: '" + std::string(b, e) + "'");
        }
        push(kQasmCx, qs[0], qs[1], pv, 0);
      } else if (known && kit->second == kQasmSwap && nArgs == 2) {
        // only in mapped circuits; the mapper itself wants SWAPs decomposed
        push(kQasmSwap, qs[0], qs[1], pv, 0);
      } else if (known && kit->second != kQasmSwap) {
        if (nArgs != 1) {
          throw std::invalid_argument("bad 1q gate: '" + std::string(b, e) + "'");
        }
        push(kit->second, -1, qs[0], pv, np);
      } else {
        // Mapper::createLayers :157-161: gates with more than one control
        throw std::invalid_argument(
            "Circuit contains gates with more than one control or an "
            "unsupported gate (" + head + "). Please make sure that the input "
            "circuit's gates are decomposed to the appropriate gate set!");
      }
    }
  }
  void push(uint8_t k, int32_t c, int32_t t, const double* pv, uint8_t np) {
    kind.push_back(k);
    control.push_back(c);
    target.push_back(t);
    params.insert(params.end(), pv, pv + 3);
    nParams.push_back(np);
  }

  // Mapper's view (qmap_b200/circuit.py::mapper_view): operand q becomes the
  // rank of q among the qubits that are not idle (touched by a gate or
  // measured); identity[p] = p for every original used index p < nPhysical
  void mapperView(uint32_t nPhysical, int32_t* c, int32_t* t,
                  int16_t* identity) const {
    for (const uint8_t k : kind) {
      if (k == kQasmSwap) {
        throw std::invalid_argument(
            "Circuit contains SWAP gates. Please make sure that the input "
            "circuit's gates are decomposed to the appropriate gate set!");
      }
    }
    std::vector<int32_t> lut(std::max<uint32_t>(nQubits, 1), -1);
    std::vector<uint8_t> used(std::max<uint32_t>(nQubits, 1), 0);
    for (size_t i = 0; i < kind.size(); ++i) {
      used[static_cast<size_t>(target[i])] = 1;
      if (control[i] >= 0) {
        used[static_cast<size_t>(control[i])] = 1;
      }
    }
    for (const uint32_t q : measQubit) {
      used[q] = 1;
    }
    int32_t r = 0;
    for (uint32_t q = 0; q < nQubits; ++q) {
      if (used[q] != 0) {
        lut[q] = r++;
      }
    }
    for (size_t i = 0; i < kind.size(); ++i) {
      t[i] = lut[static_cast<size_t>(target[i])];
      c[i] = control[i] >= 0 ? lut[static_cast<size_t>(control[i])] : -1;
    }
    if (identity != nullptr) {
      for (uint32_t p = 0; p < nPhysical; ++p) {
        identity[p] = (p < nQubits && used[p] != 0) ? static_cast<int16_t>(p)
                                                    : static_cast<int16_t>(-1);
      }
    }
  }
};

}  // namespace qmg
