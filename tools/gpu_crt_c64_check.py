"""C64 (BASELINE configs[1]): the two forms of the int8 metric product side by side in one process --
digit planes (default) and residues (ISDF_SYRK_CRT=16): W, J and K differences.  The knob is read on every
call of isdf_syrk_i8, so it can be switched between two builds."""
import json
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import workloads
from pyscf_isdf_b200.isdf import ISDF

cell = workloads.make_cell(sys.argv[1] if len(sys.argv) > 1 else 'c64')
dm = workloads.synthetic_dm(cell.nao)
out = {}
res = {}
for name, env in (('planes', None), ('crt16', '16')):
    if env is None:
        os.environ.pop('ISDF_SYRK_CRT', None)
    else:
        os.environ['ISDF_SYRK_CRT'] = env
    mydf = ISDF(cell, c=15, keep_aoR=False).build()
    vj, vk = mydf.get_jk(dm)
    res[name] = (mydf.W.clone(), vj, vk, np.array(mydf.ip_index))
    out[name + '_wgemm_ms'] = mydf.timings.get('wgemm_ms')
    del mydf
W0, vj0, vk0, ip0 = res['planes']
W1, vj1, vk1, ip1 = res['crt16']
out['same_points'] = bool(np.array_equal(ip0, ip1))
out['W_max_abs'] = float(W0.abs().max())
out['W_diff_max'] = float((W0 - W1).abs().max())
out['vj_diff_max'] = float(abs(vj0 - vj1).max())
out['vk_diff_max'] = float(abs(vk0 - vk1).max())
out['vj_max'] = float(abs(vj0).max())
out['vk_max'] = float(abs(vk0).max())
print(json.dumps(out))
