import sys, time, numpy as np
sys.path.insert(0, '.')
import openflash_b200 as ofb
geo = ofb.BasicRegionGeometry.from_vectors(a=np.array([0.125, 0.25]), d=np.array([0.25, 0.125]), h=1.001, NMK=[10] * 3, body_map=[0, 0], heaving_map=[True])
prob = ofb.MEEMProblem(geo)
prob.set_frequencies(np.array([float(np.sqrt(np.tanh(1.001) * 9.81))]))
eng = ofb.MEEMEngine([prob], device=0)
lat = []
for _ in range(45):
    t0 = time.perf_counter(); ds = eng.run_and_store_results(0).dataset; lat.append(1e3 * (time.perf_counter() - t0))
print("latency ms median", round(float(np.median(lat[5:])), 4), "min", round(min(lat), 4))
