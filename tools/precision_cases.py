import sys,time,json,os
sys.path.insert(0,".")
import numpy as np
import qflex_b200
from tests import support
lib=qflex_b200.default_library()
gold=json.load(open("tests/golden/amplitudes.json"))
for name in ["sycamore_53_10_1slice","rect7x7_1-32-1_4amps","sycamore_53_8_64amps"]:
    c,o,g=support.case_texts(support.CASES[name])
    want=np.array(gold[name]["re"])+1j*np.array(gold[name]["im"])
    for prec in ("ffma","auto"):
        res,st=lib.evaluate(c,o,g,memory_limit=100<<30,precision=prec)
        print(name,prec,os.environ.get("QFLEX_TC_CHUNK"),"err=%.2e"%support.rel_err([a for _,a in res],want),"ms=%.1f"%st["device_ms_contract"],flush=True)
