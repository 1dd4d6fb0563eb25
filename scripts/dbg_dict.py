import os, sys
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import scenes
import titan_b200 as tb

def mk(sc, dict_on):
    os.environ["TB_TPM"] = "1"; os.environ["TB_DICT"] = dict_on
    e = tb.Engine(0, tb.MATH_FAST); scenes.upload(e, sc); e.run(60, 0.0)
    return e

sc = scenes.c2_lattice(24)
t = 0.0
for _ in range(60): t += sc.dt
for label, sel in (("few", slice(5, 40, 3)), ("many", slice(0, None, 5))):
    plain, dic = mk(sc, "0"), mk(sc, "1")
    k2 = sc.k.copy(); k2[sel] *= 1.5
    # fresh engines that start from the 60-step state with k2 already in place
    st = plain.get_mass_state()
    sc2 = sc.copy(); sc2.pos[:] = st["pos"]; sc2.vel[:] = st["vel"]; sc2.k[:] = k2
    res = {}
    for name, e in (("plain", plain), ("dict", dic)):
        e.set_spring_params(k=k2); e.run(50, t); res[name] = e.get_mass_state(); print(label, name, e.info()["dict_records"])
    for name, d in (("fresh_plain", "0"), ("fresh_dict", "1")):
        os.environ["TB_TPM"] = "1"; os.environ["TB_DICT"] = d
        e = tb.Engine(0, tb.MATH_FAST); scenes.upload(e, sc2); e.run(50, t); res[name] = e.get_mass_state(); print(label, name, e.info()["dict_records"]); e.close()
    names = list(res)
    for i in range(len(names)):
        for j in range(i + 1, len(names)):
            d = np.abs(res[names[i]]["pos"] - res[names[j]]["pos"])
            print(label, names[i], "vs", names[j], "max abs", d.max(), "n differing masses", int((d.max(axis=1) > 0).sum()))
    plain.close(); dic.close()
