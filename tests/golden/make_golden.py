"""Transcribes the reference's text known-answer vectors into golden.json (run once in the build container;
reads only /root/reference text files to cross-check the transcription where possible)."""
import json
import os
import re

G = {
    "csr6": {  # ut_petsc.f90:83-149 (both ranks' rows concatenated)
        "m": 6, "n": 6, "ptr": [1, 2, 4, 6, 8, 9, 11], "ind": [1, 1, 2, 2, 4, 4, 5, 5, 5, 6],
        "val": [1, 2, 3, 4, 5, 6, 7, 8, 9, 10], "x": [1, 2, 3, 4, 5, 6], "y": [1, 8, 28, 59, 40, 105]},
    "csr7_1": {  # ut_mpi_CSR_matrix_vector_multiplication.f90:186-246
        "m": 7, "n": 7, "ptr": [1, 2, 4, 6, 9, 11, 13, 15], "ind": [1, 2, 3, 2, 5, 3, 4, 6, 4, 5, 5, 7, 6, 7],
        "val": [1, 2, 3, 4, 1, 2, 3, 4, 1, 2, 3, 4, 1, 2], "x": [1, 2, 3, 4, 1, 2, 3], "y": [1, 13, 9, 26, 6, 15, 8]},
    "csr7_2": {  # ut_mpi_CSR_matrix_vector_multiplication.f90:248-318
        "m": 7, "n": 7, "ptr": [1, 3, 6, 8, 13, 16, 19, 22],
        "ind": [1, 5, 2, 3, 6, 2, 5, 1, 3, 4, 6, 7, 2, 4, 5, 1, 5, 7, 2, 6, 7],
        "val": [1, 5, 2, 3, 6, 4, 1, 5, 2, 3, 4, 5, 6, 1, 2, 5, 3, 4, 6, 1, 2], "x": [1, 2, 3, 4, 1, 2, 3],
        "y": [6, 25, 9, 46, 18, 20, 20]},
    "partition_100_4": {"vi1": [1, 26, 51, 76], "vi2": [25, 50, 75, 100]},  # mesh_parallelisation.f90:3-36
    "dummy5": {  # mesh_dummy_meshes.f90:67-100
        "C": [[2, 5, 4, 0], [3, 5, 1, 0], [4, 5, 2, 0], [1, 5, 3, 0], [1, 2, 3, 4]],
        "iTri": [[1, 4, 0, 0], [2, 1, 0, 0], [3, 2, 0, 0], [4, 3, 0, 0], [1, 2, 3, 4]],
        "TriC": [[2, 4, 0], [3, 1, 0], [4, 2, 0], [1, 3, 0]], "nC": [3, 3, 3, 3, 4], "niTri": [2, 2, 2, 2, 4],
        "VBI": [6, 4, 2, 8, 0]},
    "mismip_mod_small": {"nV": 1022},
}

ref = "/root/reference/automated_testing/UFEMISM/integrated_test_MISMIP_mod_small/reference/checksum_logfile.txt"
if os.path.exists(ref):
    txt = open(ref).read().splitlines()
    G["mismip_mod_small"]["petsc_version_line"] = next((l.strip() for l in txt[:12] if "PETSc" in l), "")
    for l in txt[:200]:
        mm = re.match(r"\s*(mask_grounded_ice|mask_floating_ice|mask_icefree_land|mask_icefree_ocean)\s*:\s*(.*)", l)
        if mm and mm.group(1) not in G["mismip_mod_small"]:
            G["mismip_mod_small"][mm.group(1)] = mm.group(2).strip()
    # output of the reference itself: (sum, sum|.|, min, max) of u_b after each of the first 20 relaxed DIVA viscosity iterations of the first solve
    rows = [l for l in txt if l.startswith(" u_b:") and "/solve_DIVA_infinite_slab/relax_viscosity_iterations]" in l][:20]
    G["mismip_mod_small"]["u_b_relaxed"] = [[float(x.replace("-.", "-0.")) for x in re.findall(r"=\s*(-?[0-9]*\.[0-9]+E[+-][0-9]+)", l)] for l in rows]
    # the first solve as a whole: number of viscosity iterations until convergence, the converged DIVA%u_vav_b and the driving stress
    first = next(i for i, l in enumerate(txt) if l.startswith(" DIVA%u_vav_b:"))
    num = lambda l: [float(x.replace("-.", "-0.")) for x in re.findall(r"=\s*(-?[0-9]*\.[0-9]+E[+-][0-9]+)", l)]
    G["mismip_mod_small"]["n_visc_its_first_solve"] = sum(1 for l in txt[:first] if l.startswith(" u_b:") and "relax_viscosity_iterations]" in l)
    G["mismip_mod_small"]["u_vav_b_converged"] = num(txt[first])
    G["mismip_mod_small"]["tau_dx_b"] = num(next(l for l in txt if l.startswith(" tau_dx_b:")))
    # the first thickness step that sees velocities (dt = 0.1 yr, right after that solve): flux divergence of the uniform 100 m slab
    G["mismip_mod_small"]["divQ_after_first_solve"] = num(next(l for l in txt[first:] if l.startswith(" divQ:")))
    # the whole 100-year run: per predictor-corrector iteration the number of viscosity iterations of its stress-balance solve, the converged
    # DIVA%u_vav_b, the corrected thickness and the truncation error; first and last grounded fractions and mask counts
    T = {"n_visc_its": [], "u_vav_b": [], "Hi_np1": [], "tau_np1": [], "fraction_gr_b": [], "fraction_gr": [], "mask_grounded_ice": [], "mask_floating_ice": []}
    cnt = 0
    for l in txt:
        if l.startswith(" u_b:") and "relax_viscosity_iterations]" in l:
            cnt += 1
        elif l.startswith(" DIVA%u_vav_b:"):
            T["n_visc_its"].append(cnt); cnt = 0; T["u_vav_b"].append(num(l))
        elif l.startswith(" region%ice%pc%Hi_np1:"):
            T["Hi_np1"].append(num(l))
        elif l.startswith(" pc%tau_np1:"):
            T["tau_np1"].append(num(l))
        elif l.startswith(" fraction_gr_b:") or l.startswith(" fraction_gr:"):
            T[l.split(":")[0].strip()].append(num(l)[0])
        elif l.startswith(" mask_grounded_ice:") or l.startswith(" mask_floating_ice:"):
            T[l.split(":")[0].strip()].append(int(re.search(r"sum = (\d+)", l).group(1)))
    for k in ("fraction_gr_b", "fraction_gr", "mask_grounded_ice", "mask_floating_ice"):
        T[k] = [T[k][0], T[k][-1]]
    G["mismip_mod_small"]["time_loop"] = T

# Output of the reference itself: the checksum log of integrated_test_SSA_icestream_small (16 km, SSA, BC 'infinite_SSA_icestream').  Its first
# stress-balance solve logs (sum, sum|.|, min, max) of u_b / v_b after every relaxed viscosity iteration: 520 iterations, all kept.
def _nums(line):
    return [float(x.replace("-.", "-0.")) for x in re.findall(r"=\s*(-?[0-9]*\.[0-9]+E[+-][0-9]+)", line)]


ref = "/root/reference/automated_testing/UFEMISM/integrated_test_SSA_icestream_small/reference/checksum_logfile.txt"
if os.path.exists(ref):
    S = {"source": "automated_testing/UFEMISM/integrated_test_SSA_icestream_small/reference/checksum_logfile.txt",
         "columns": ["sum", "sum_abs", "min", "max"], "u_b_relaxed": [], "v_b_relaxed": [], "u_b_limited": []}
    for l in open(ref):
        if l.startswith(" mask_grounded_ice:") and "nV" not in S:
            S["nV"] = int(re.search(r"sum = (\d+)", l).group(1))          # every vertex is grounded ice in this experiment
        elif l.startswith(" fraction_gr_b:") and "nTri" not in S:
            S["nTri"] = int(round(_nums(l)[0]))                            # fraction_gr_b = 1 on every triangle
        elif l.startswith(" tau_dx_b:") and "tau_dx_b" not in S:
            S["tau_dx_b"] = _nums(l)
        elif l.startswith(" ice%Hi:") and "Hi" not in S:
            S["Hi"] = _nums(l)
        elif "/solve_SSA/relax_viscosity_iterations]" in l and l.startswith(" u_b:"):
            S["u_b_relaxed"].append(_nums(l))
        elif "/solve_SSA/relax_viscosity_iterations]" in l and l.startswith(" v_b:"):
            S["v_b_relaxed"].append(_nums(l))
        elif "/solve_SSA/apply_velocity_limits]" in l and l.startswith(" u_b:"):
            S["u_b_limited"].append(_nums(l))
    json.dump(S, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ssa_icestream_small.json"), "w"))   # compact: 520 x 3 rows

# Output of the reference itself: ISMIP-HOM experiment A, L = 160 / 80 / 40 km, DIVA, no sliding, 'periodic_ISMIP-HOM' borders (integrated_test_ISMIP_HOM_small):
# u_b after each of the 48-50 relaxed viscosity iterations, and the statistics of every named DIVA field the last iteration logged.
for _L in (160, 80, 40):
    ref = "/root/reference/automated_testing/UFEMISM/integrated_test_ISMIP_HOM_small/reference/checksum_logfile_A_%d_DIVA.txt" % _L
    if not os.path.exists(ref):
        continue
    S = {"source": "automated_testing/UFEMISM/integrated_test_ISMIP_HOM_small/reference/checksum_logfile_A_%d_DIVA.txt" % _L,
         "columns": ["sum", "sum_abs", "min", "max"], "u_b_relaxed": [], "last": {}}
    for l in open(ref):
        if l.startswith(" mask_grounded_ice:") and "nV" not in S:
            S["nV"] = int(re.search(r"sum = (\d+)", l).group(1))
        elif l.startswith(" fraction_gr_b:") and "nTri" not in S:
            S["nTri"] = int(round(_nums(l)[0]))
        elif "/solve_DIVA_infinite_slab" in l:
            name = l.split(":")[0].strip()
            if name == "u_b" and "relax_viscosity_iterations]" in l:
                S["u_b_relaxed"].append(_nums(l))
            if not (name in ("u_b", "v_b") and "apply_velocity_limits]" in l):
                S["last"][name] = _nums(l)          # later lines overwrite earlier ones: the last viscosity iteration stays
    G["ismip_hom_A_%d_DIVA" % _L] = S

# Output of the reference itself: the Halfar dome run of integrated_test_Halfar_dome_small (100 yr, SIA, predictor-corrector time stepping).  Every
# predictor-corrector iteration logs the corrected thickness and the truncation error: 58 iterations, all kept.  (A three-digit exponent is printed without
# its E: 0.76841014953792-234.)
def _nums3(line):
    return [float(re.sub(r"(\d)([+-]\d{3})$", r"\1E\2", x).replace("-.", "-0.")) for x in re.findall(r"=\s*(-?[0-9]*\.[0-9]+(?:E[+-][0-9]+|[+-][0-9]{3}))", line)]


ref = "/root/reference/automated_testing/UFEMISM/integrated_test_Halfar_dome_small/reference/checksum_logfile_Halfar_40km_grid.txt"
if os.path.exists(ref):
    S = {"source": "automated_testing/UFEMISM/integrated_test_Halfar_dome_small/reference/checksum_logfile_Halfar_40km_grid.txt",
         "columns": ["sum", "sum_abs", "min", "max"], "Hi_np1": [], "Hi_star_np1": [], "tau_np1": [], "divQ": []}
    for l in open(ref):
        if l.startswith(" fraction_gr:") and "nV" not in S:
            S["nV"] = int(round(_nums3(l)[0]))                              # fraction_gr = 1 on every vertex
        elif l.startswith(" fraction_gr_b:") and "nTri" not in S:
            S["nTri"] = int(round(_nums3(l)[0]))
        elif l.startswith(" ice%Hi:") and "Hi_initial" not in S:
            S["Hi_initial"] = _nums3(l)
        elif l.startswith(" region%ice%pc%Hi_np1:"):
            S["Hi_np1"].append(_nums3(l))
        elif l.startswith(" region%ice%pc%Hi_star_np1:"):
            S["Hi_star_np1"].append(_nums3(l))
        elif l.startswith(" pc%tau_np1:"):
            S["tau_np1"].append(_nums3(l))
        elif l.startswith(" divQ:"):
            S["divQ"].append(_nums3(l))
        elif l.startswith(" region%ice%Hs:"):
            S["Hs_final"] = _nums3(l)
    json.dump(S, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "halfar_40km.json"), "w"))

json.dump(G, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json"), "w"), indent=1)
print("wrote golden.json")
