// Stand-in for mqt-core's qasm3::Importer, restricted to the OpenQASM 2
// subset the reference's examples/*.qasm use (qreg/creg, cx, h, t, tdg, x, y,
// z, s, sdg, rz, u1, u2, u3, measure, barrier). TEST INFRASTRUCTURE.
#pragma once
#include "ir/QuantumComputation.hpp"
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>

namespace qasm3 {
class Importer {
  struct Expr {
    const std::string& s;
    std::size_t p = 0;
    void ws() {
      while (p < s.size() && std::isspace(static_cast<unsigned char>(s[p]))) {
        ++p;
      }
    }
    double atom() {
      ws();
      if (p < s.size() && s[p] == '(') {
        ++p;
        const double v = sum();
        ws();
        ++p;
        return v;
      }
      if (p < s.size() && (s[p] == '-' || s[p] == '+')) {
        const bool neg = s[p] == '-';
        ++p;
        const double v = atom();
        return neg ? -v : v;
      }
      if (s.compare(p, 2, "pi") == 0) {
        p += 2;
        return 3.14159265358979323846;
      }
      char* e = nullptr;
      const double v = std::strtod(s.c_str() + p, &e);
      p = static_cast<std::size_t>(e - s.c_str());
      return v;
    }
    double prod() {
      double v = atom();
      for (ws(); p < s.size() && (s[p] == '*' || s[p] == '/'); ws()) {
        const char o = s[p++];
        const double r = atom();
        v = o == '*' ? v * r : v / r;
      }
      return v;
    }
    double sum() {
      double v = prod();
      for (ws(); p < s.size() && (s[p] == '+' || s[p] == '-'); ws()) {
        const char o = s[p++];
        const double r = prod();
        v = o == '+' ? v + r : v - r;
      }
      return v;
    }
  };

public:
  static qc::QuantumComputation imports(const std::string& text) {
    std::istringstream is(text);
    return import(is);
  }
  static qc::QuantumComputation importf(const std::string& filename) {
    std::ifstream ifs(filename);
    if (!ifs.good()) {
      throw std::runtime_error("Could not open file " + filename);
    }
    auto qc = import(ifs);
    const auto slash = filename.find_last_of('/');
    const auto dot = filename.find_last_of('.');
    qc.setName(filename.substr(slash + 1, dot - slash - 1));
    return qc;
  }
  static qc::QuantumComputation import(std::istream& is) {
    qc::QuantumComputation circ;
    std::map<std::string, qc::Qubit> qregStart;
    std::map<std::string, std::size_t> qregSize;
    std::map<std::string, qc::Bit> cregStart;
    std::size_t nq = 0;
    std::size_t nc = 0;
    std::string stmt;
    while (std::getline(is, stmt, ';')) {
      // strip comments and leading whitespace
      std::string t;
      {
        std::istringstream ls(stmt);
        std::string line;
        while (std::getline(ls, line)) {
          const auto c = line.find("//");
          if (c != std::string::npos) {
            line = line.substr(0, c);
          }
          t += line + " ";
        }
      }
      const auto b = t.find_first_not_of(" \t\r\n");
      if (b == std::string::npos) {
        continue;
      }
      t = t.substr(b);
      auto head = t.substr(0, t.find_first_of(" \t(["));
      if (head == "OPENQASM" || head == "include") {
        continue;
      }
      auto parseReg = [&](const std::string& tok, std::string& nm,
                          long& idx) {
        const auto lb = tok.find('[');
        if (lb == std::string::npos) {
          nm = tok;
          idx = -1;
        } else {
          nm = tok.substr(0, lb);
          idx = std::stol(tok.substr(lb + 1));
        }
        const auto f = nm.find_first_not_of(" \t");
        const auto l = nm.find_last_not_of(" \t");
        nm = nm.substr(f, l - f + 1);
      };
      if (head == "qreg" || head == "creg") {
        std::string nm;
        long sz = 0;
        parseReg(t.substr(4), nm, sz);
        if (head == "qreg") {
          qregStart[nm] = static_cast<qc::Qubit>(nq);
          qregSize[nm] = static_cast<std::size_t>(sz);
          circ.addQubitRegister(static_cast<std::size_t>(sz), nm);
          nq += static_cast<std::size_t>(sz);
        } else {
          cregStart[nm] = nc;
          circ.addClassicalRegister(static_cast<std::size_t>(sz), nm);
          nc += static_cast<std::size_t>(sz);
        }
        continue;
      }
      // gate [ (params) ] args
      std::vector<double> params;
      std::string rest = t.substr(head.size());
      {
        const auto f = rest.find_first_not_of(" \t");
        if (f != std::string::npos && rest[f] == '(') {
          int depth = 0;
          std::size_t e = f;
          for (; e < rest.size(); ++e) {
            if (rest[e] == '(') {
              ++depth;
            } else if (rest[e] == ')' && --depth == 0) {
              break;
            }
          }
          std::string ps = rest.substr(f + 1, e - f - 1);
          rest = rest.substr(e + 1);
          std::size_t start = 0;
          int d = 0;
          for (std::size_t i = 0; i <= ps.size(); ++i) {
            if (i == ps.size() || (ps[i] == ',' && d == 0)) {
              const std::string one = ps.substr(start, i - start);
              Expr ex{one};
              params.push_back(ex.sum());
              start = i + 1;
            } else if (ps[i] == '(') {
              ++d;
            } else if (ps[i] == ')') {
              --d;
            }
          }
        }
      }
      std::vector<qc::Qubit> qs;
      std::vector<std::string> argNames;
      if (head == "measure") {
        const auto arrow = rest.find("->");
        std::string nm;
        long idx = 0;
        parseReg(rest.substr(0, arrow), nm, idx);
        std::string cn;
        long cidx = 0;
        parseReg(rest.substr(arrow + 2), cn, cidx);
        if (idx < 0) {
          for (std::size_t i = 0; i < qregSize.at(nm); ++i) {
            circ.measure(qregStart.at(nm) + static_cast<qc::Qubit>(i),
                         cregStart.at(cn) + i);
          }
        } else {
          circ.measure(qregStart.at(nm) + static_cast<qc::Qubit>(idx),
                       cregStart.at(cn) + static_cast<qc::Bit>(cidx));
        }
        continue;
      }
      {
        std::istringstream as(rest);
        std::string tok;
        while (std::getline(as, tok, ',')) {
          std::string nm;
          long idx = 0;
          parseReg(tok, nm, idx);
          if (idx < 0) {
            if (head == "barrier") {
              for (std::size_t i = 0; i < qregSize.at(nm); ++i) {
                qs.push_back(qregStart.at(nm) + static_cast<qc::Qubit>(i));
              }
              continue;
            }
            throw std::runtime_error("register broadcast unsupported: " + t);
          }
          qs.push_back(qregStart.at(nm) + static_cast<qc::Qubit>(idx));
        }
      }
      using namespace qc;
      if (head == "barrier") {
        circ.emplace_back<StandardOperation>(qs, Barrier);
      } else if (head == "cx" || head == "CX") {
        circ.cx(Control{qs.at(0)}, qs.at(1));
      } else if (head == "swap") {
        circ.swap(qs.at(0), qs.at(1));
      } else if (head == "ccx") {
        circ.emplace_back<StandardOperation>(
            Controls{Control{qs.at(0)}, Control{qs.at(1)}}, qs.at(2), X);
      } else {
        OpType ot = None;
        if (head == "u1" || head == "p") {
          ot = P;
        } else if (head == "u" || head == "u3" || head == "U") {
          ot = U;
        } else {
          ot = opTypeFromString(head);
        }
        circ.emplace_back<StandardOperation>(qs.at(0), ot, params);
      }
    }
    return circ;
  }
};
} // namespace qasm3
