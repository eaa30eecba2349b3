import sys, os
sys.path.insert(0,'tests')
import parity, numpy as np
from qmap_b200 import _abi, routing, architecture as ar
ra=parity.ref().arch_load(os.path.join(parity.ARCHS,'ibm_qx4.arch'))
be,a=parity.cuda_backend_for(ra)
msg,out,rr=parity.run_pair(be,_abi.load(),ra,os.path.join(parity.CIRCUITS,'4gt13_92.qasm'),lookahead_heuristic='none')
print(msg); print(out.final_qubits, out.final_locations, out.start_locations, rr.final_qubits, out.layer_qubits[-1], out.placements, out.status, out.n_layers_done)
