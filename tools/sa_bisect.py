import subprocess, sys, os
code = '''
import ctypes as C, sys, torch
sys.path.insert(0, "%s")
from zipvoice_b200 import _lib
lib = _lib.load(); torch.zeros(1, device="cuda")
ms = C.c_float()
rc = lib.zv_debug_attention_bench(4, 300, 4, 4, 12, 0, int(sys.argv[1]), 1, C.byref(ms))
torch.cuda.synchronize()
print("dbg", sys.argv[1], "rc", rc, "ms", ms.value)
''' % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for dbg in (15, 14, 13, 11, 7, 0):
    r = subprocess.run([sys.executable, "-c", code, str(dbg)], capture_output=True, text=True, timeout=120)
    print(r.stdout.strip()[-200:], "|", r.stderr.strip()[-300:].replace("\n", " "))
