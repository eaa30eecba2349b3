/* route_qasm.c - the C ABI of libqmap_b200.so from plain C: read an OpenQASM 2
 * file and a coupling map (.arch: qubit count, then one "a b" edge per line),
 * route the circuit with the reference's defaults and print the SWAPs the
 * router inserts before every layer.
 *
 *   gcc -std=c11 -Iinclude examples/route_qasm.c -Lmqt-qmap_b200/csrc \
 *       -lqmap_b200 -Wl,-rpath,$PWD/mqt-qmap_b200/csrc -o route_qasm
 *   ./route_qasm circuit.qasm device.arch
 *
 * What each call replaces in the reference is written in include/qmap_b200.h.
 * Exit code 3: no CUDA device (the library has no CPU fallback).
 */
#include <stdio.h>
#include <stdlib.h>

#include "qmap_b200.h"

static void die(const char* what) {
  fprintf(stderr, "%s: %s\n", what, qmg_last_error());
  exit(1);
}

int main(int argc, char** argv) {
  if (argc != 3) {
    fprintf(stderr, "usage: %s circuit.qasm device.arch\n", argv[0]);
    return 2;
  }
  /* ---- the coupling map (Architecture::loadCouplingMap, Architecture.cpp:52-83) */
  FILE* f = fopen(argv[2], "r");
  if (f == NULL) {
    fprintf(stderr, "cannot open %s\n", argv[2]);
    return 2;
  }
  unsigned n = 0, a = 0, b = 0;
  if (fscanf(f, "%u", &n) != 1 || n == 0 || n > 4096) {
    fprintf(stderr, "no qubit count in %s\n", argv[2]);
    return 2;
  }
  size_t nEdges = 0, cap = 64;
  uint16_t* edges = malloc(2 * cap * sizeof(uint16_t));
  while (fscanf(f, "%u %u", &a, &b) == 2) {
    if (nEdges == cap) {
      cap *= 2;
      edges = realloc(edges, 2 * cap * sizeof(uint16_t));
    }
    edges[2 * nEdges] = (uint16_t)a;
    edges[2 * nEdges + 1] = (uint16_t)b;
    ++nEdges;
  }
  fclose(f);

  /* ---- the circuit as Mapper sees it (Mapper.cpp:33-53, 166-172) */
  qmg_qasm* qasm = NULL;
  if (qmg_qasm_read_file(argv[1], &qasm) != QMG_OK) {
    die("reading the circuit");
  }
  qmg_qasm_info info;
  qmg_qasm_get_info(qasm, &info);
  int32_t* control = malloc((info.n_gates + 1) * sizeof(int32_t));
  int32_t* target = malloc((info.n_gates + 1) * sizeof(int32_t));
  if (qmg_qasm_mapper_view(qasm, n, control, target, NULL) != QMG_OK) {
    die("the mapper's view of the circuit");
  }

  /* ---- Configuration defaults, createLayers */
  qmg_config cfg;
  qmg_config_default(&cfg);
  qmg_circuit* circuit = NULL;
  if (qmg_circuit_create(n, (size_t)info.n_gates, control, target, cfg.layering, &circuit) != QMG_OK) {
    die("createLayers");
  }
  qmg_instance inst = {0};
  qmg_circuit_layers(circuit, &inst.layers);
  const uint32_t nLayers = inst.layers.n_layers;
  /* room for the layers automatic splits can add (the library then splits
   * layers itself, like Mapper::splitLayer + the restart in aStarMap) */
  inst.layer_capacity = nLayers + (uint32_t)info.n_gates;

  /* ---- the device-resident Architecture with its distance tables */
  if (qmg_device_count() < 1) {
    fprintf(stderr, "no CUDA device: libqmap_b200 has no CPU fallback\n");
    return 3;
  }
  qmg_arch* arch = NULL;
  if (qmg_arch_create((uint16_t)n, edges, nEdges, 0, &arch) != QMG_OK) {
    die("qmg_arch_create");
  }

  /* ---- route: one launch for the whole circuit */
  qmg_route_result res = {0};
  const uint32_t layerCap = inst.layer_capacity;
  res.swap_capacity = 64 * layerCap + 1024;
  res.swaps = malloc(2 * (size_t)res.swap_capacity * sizeof(uint16_t));
  res.layer_n_swaps = calloc(layerCap, sizeof(uint16_t));
  res.final_qubits = malloc(n * sizeof(int16_t));
  res.final_locations = malloc(n * sizeof(int16_t));
  res.start_locations = malloc(n * sizeof(int16_t));
  res.placements = malloc(n * sizeof(qmg_placement));
  res.placement_capacity = n;
  res.split_layers = malloc(((size_t)info.n_gates + 1) * sizeof(uint32_t));
  res.split_capacity = (uint32_t)info.n_gates;
  qmg_route_stats stats;
  if (qmg_route_batch(arch, &cfg, 1, &inst, &res, NULL, &stats) != QMG_OK || res.status != QMG_OK) {
    fprintf(stderr, "routing stopped with status %d: %s\n", (int)res.status, qmg_last_error());
    return 1;
  }

  printf("%u qubits, %llu gates, %u layers (%u after %u splits); %llu nodes expanded, kernel %.3f ms\n",
         info.n_qubits, (unsigned long long)info.n_gates, nLayers, res.n_layers_done, res.n_splits,
         (unsigned long long)res.expanded_nodes, stats.kernel_ms);
  size_t k = 0;
  for (uint32_t l = 0; l < res.n_layers_done; ++l) {
    for (uint16_t s = 0; s < res.layer_n_swaps[l]; ++s, ++k) {
      printf("layer %u: SWAP %u %u\n", l, res.swaps[2 * k], res.swaps[2 * k + 1]);
    }
  }
  printf("final layout (physical <- logical):");
  for (unsigned p = 0; p < n; ++p) {
    printf(" %d", (int)res.final_qubits[p]);
  }
  printf("\n");

  qmg_arch_destroy(arch);
  qmg_circuit_destroy(circuit);
  qmg_qasm_destroy(qasm);
  return 0;
}
