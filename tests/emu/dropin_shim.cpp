// dropin_shim.cpp - TEST INFRASTRUCTURE ONLY.
//
// Lets the compiled C++ drop-in (integration/GpuHeuristicMapper.hpp around the
// reference's own sources) run in a container without a GPU: the five entry
// points of include/qmap_b200.h the drop-in calls are answered by the host
// debug build of the kernel source (emu.cpp).  Cost tables are not computed
// here (the table kernels exist only in CUDA): they are taken from the
// reference Architecture the test driver registers before map().
// Built into oracle/_ref/libqmapdropin_emu.so by `make -C oracle dropin_emu`;
// loaded only by tests/test_dropin_host.py; never part of the product.
#include "emu.cpp"

#include "sc/Architecture.hpp"

struct qmg_arch {
  qemu_arch* emu = nullptr;
};

namespace {
const Architecture* gReference = nullptr;
std::vector<double> flatMatrix(const Matrix& m) {
  std::vector<double> v;
  for (const auto& row : m) {
    v.insert(v.end(), row.begin(), row.end());
  }
  return v;
}
}  // namespace

// the Architecture whose tables the next qmg_arch_create takes
QMG_API void qemu_shim_reference_architecture(const void* architecture) {
  gReference = static_cast<const Architecture*>(architecture);
}

QMG_API int qmg_arch_create(uint16_t n, const uint16_t* edges, size_t nEdges,
                            int /*device*/, qmg_arch** out) {
  if (gReference == nullptr || gReference->getNqubits() != n) {
    return fail(QMG_BAD_ARG, "dropin_shim: no reference architecture registered");
  }
  const auto d0 = flatMatrix(gReference->getDistanceTable(false));
  const auto d1 = flatMatrix(gReference->getDistanceTable(true));
  auto a = std::make_unique<qmg_arch>();
  const int rc = qemu_arch_create(n, edges, nEdges, d0.data(), d1.data(), &a->emu);
  if (rc != QMG_OK) {
    return rc;
  }
  *out = a.release();
  return QMG_OK;
}

QMG_API int qmg_arch_set_fidelity_costs(qmg_arch* arch, const double* c1,
                                        const double* c2, const double* cswap) {
  if (gReference == nullptr || !gReference->isFidelityAvailable()) {
    return fail(QMG_BAD_ARG, "dropin_shim: the reference architecture has no calibration");
  }
  std::vector<double> fid;
  const auto& tables = gReference->getFidelityDistanceTables();
  for (const auto& t : tables) {
    const auto f = flatMatrix(t);
    fid.insert(fid.end(), f.begin(), f.end());
  }
  return qemu_arch_set_fidelity(arch->emu, c1, c2, cswap, fid.data(),
                                static_cast<uint32_t>(tables.size()));
}

QMG_API void qmg_arch_destroy(qmg_arch* arch) {
  if (arch != nullptr) {
    qemu_arch_destroy(arch->emu);
    delete arch;
  }
}

QMG_API int qmg_route_batch(const qmg_arch* arch, const qmg_config* cfg,
                            size_t nInstances, const qmg_instance* instances,
                            qmg_route_result* results,
                            const qmg_route_options* /*options*/,
                            qmg_route_stats* stats) {
  if (stats != nullptr) {
    *stats = qmg_route_stats{};
  }
  return qemu_route_batch(arch->emu, cfg, nInstances, instances, results, 0);
}
