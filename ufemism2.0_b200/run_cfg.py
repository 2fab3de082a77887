"""Run the stress balance of one of the reference's idealised experiments from its .cfg file through the C ABI, and print the result the way the
reference's checksum log does -- the .cfg-driven, switch-selected use of this path that INTEGRATION.md describes, with the ctypes driver standing in for
the Fortran host (no Fortran compiler in this image).

    python ufemism2.0_b200/run_cfg.py <file.cfg> [--resolution METRES] [--backend B200] [--device 0]

What is taken from the .cfg (cfg_namelist.py): the domain (`xmin/xmax/ymin/ymax_ANT_config`), the idealised geometry
(`choice_refgeo_init_idealised_config`: MISMIP_mod, SSA_icestream, ISMIP-HOM_A, Halfar), the resolution (`maximum_resolution_uniform_config`), the stress-balance
approximation and every C% scalar of ufe_b200_config.  The mesh is a uniform mesh of this repo's generator (the reference's mesh generator is out of scope),
so the numbers are comparable with a reference log statistically, not vertex by vertex (tests/golden/README.md).

`choice_linear_algebra_backend_config = 'B200'` (or `--backend B200`) selects this library; with 'PETSc' the reference's own path is meant, which this
repository does not contain -- the script then stops, it never falls back to a CPU solve.
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cfg_namelist as cn  # noqa: E402
import ufe_b200 as ub  # noqa: E402
import workloads as wl  # noqa: E402


def fortran_e(x, digits=14):
    """a real as the reference's checksum log prints it (E format, `0.14803255706603E+02`, `-.75954187690709E+03`)"""
    if x == 0.0:
        return "0." + "0" * digits + "E+00"
    mant, exp = f"{abs(x):.{digits - 1}E}".split("E")
    s = mant.replace(".", "")
    e = int(exp) + 1
    return ("-." if x < 0 else "0.") + s + "E" + ("+" if e >= 0 else "-") + f"{abs(e):02d}"


def checksum_line(name, a, routine_path):
    """one line of checksum_logfile.txt (src/UPSY/basic/checksum_mod.f90: sum, sum(abs), min, max of a distributed field and the routine path)"""
    a = np.asarray(a, dtype=np.float64)
    return (f" {name}: sum = {fortran_e(a.sum())}, sum(abs) = {fortran_e(np.abs(a).sum())}, min = {fortran_e(a.min())}, max = {fortran_e(a.max())}"
            f" [{routine_path}]")


def workload_from_setup(setup, resolution=None, region="ANT"):
    """mesh + fields of the idealised experiment a .cfg names, with the .cfg's own configuration struct"""
    v = setup.values
    geo = v.get("choice_refgeo_init_idealised")
    xmin, xmax = float(v.get(f"xmin_{region}", -1.0e6)), float(v.get(f"xmax_{region}", 1.0e6))
    h = float(resolution if resolution else v.get("maximum_resolution_uniform", (xmax - xmin) / 40.0))
    if geo == "MISMIP_mod":
        w = wl.mismip_mod(h, half_width=0.5 * (xmax - xmin), dome=False)
        Hi0 = float(v.get("refgeo_idealised_MISMIP_mod_Hi_init", 100.0))
        w.f["Hi"] = np.where(w.f["Hi"] > 0.0, Hi0, 0.0)
        w.f.update(wl._common_fields(w.mesh, w.f["Hi"], w.f["Hb"], w.f["SL"], A_uniform=setup.cfg.uniform_Glens_flow_factor))
        if v.get("choice_basal_hydrology_model") == "Martin2011":   # calc_pore_water_pressure_Martin2011 (basal_hydrology_main.f90:227-266)
            hb_min, hb_max = float(v.get("Martin2011_hydro_Hb_min", 0.0)), float(v.get("Martin2011_hydro_Hb_max", 1000.0))
            lam = np.clip(1.0 - (w.f["Hb"] - w.f["SL"] - hb_min) / (hb_max - hb_min), 0.0, 1.0)
            w.f["effective_pressure"] = np.maximum(0.0, wl.ICE_DENSITY * wl.GRAV * w.f["Hi"] * (1.0 - 0.96 * lam))
        phi = float(v.get("slid_ZI_phi_fric_uniform", 10.0))
        w.f["till_friction_angle"] = np.full(w.mesh.nV, phi)
    elif geo == "SSA_icestream":
        w = wl.ssa_icestream(h)
    elif geo == "ISMIP-HOM_A":
        w = wl.ismip_hom_a(L=float(v["refgeo_idealised_ISMIP_HOM_L"]), h=h)
    elif geo == "Halfar":
        w = wl.halfar(h, half_width=0.5 * (xmax - xmin), H0=float(v.get("refgeo_idealised_Halfar_H0", 3000.0)),
                      R0=float(v.get("refgeo_idealised_Halfar_R0", 500e3)), A=setup.cfg.uniform_Glens_flow_factor)
    else:
        raise cn.ConfigError(f'idealised geometry "{geo}" is not one of MISMIP_mod, SSA_icestream, ISMIP-HOM_A, Halfar')
    try:
        w.f["mask_noice"] = wl.mask_noice(w.mesh, v.get("choice_mask_noice", "none"))
    except ValueError as e:
        raise cn.ConfigError(str(e)) from e
    w.cfg = setup.cfg
    w.approximation = setup.approximation
    return w


def solve_stress_balance(ctx, w, setup):
    """solve_stress_balance (conservation_of_momentum_main.f90:79-175) through the ABI: the solve, the fan-out into ice%u_3D_b etc., the secondary velocities"""
    m, cfg, L = w.mesh, w.cfg, ub.lib()
    ice = w.ice_inputs()
    st, state = ub.alloc_velocity_state(m.nV, m.nTri, m.nz)
    nv, na = C.c_int(0), C.c_int(0)
    approx = setup.approximation
    if approx in (ub.APPROX_SIA, ub.APPROX_SIASSA):
        sia = {k: np.zeros((m.nTri if k.endswith("_b") else m.nV, m.nz), order="F") for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a")}
        ub.check(L.ufe_b200_solve_SIA(ctx.h, C.byref(cfg), float(setup.values.get("SIA_maximum_diffusivity", 1e5)), ub.dp(w.f["Hi"]), ub.dp(w.f["Hs"]),
                                      ub.dp(w.f["A_flow"]), *(ub.dp(sia[k]) for k in ("u_3D_b", "v_3D_b", "D_3D_b", "du_dz_3D_a", "dv_dz_3D_a"))))
    if approx in (ub.APPROX_SSA, ub.APPROX_SIASSA):
        ub.check(L.ufe_b200_solve_SSA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
    elif approx == ub.APPROX_DIVA:
        ub.check(L.ufe_b200_solve_DIVA(ctx.h, C.byref(cfg), C.byref(ice), C.byref(st), C.byref(nv), C.byref(na), None, None, None))
    vel_s, vel = ub.alloc_ice_velocities(m.nV, m.nTri, m.nz)
    ub.check(L.ufe_b200_set_ice_velocities(ctx.h, approx, setup.hybrid_scheme, C.byref(vel_s)))
    sec_s, sec = ub.alloc_secondary(m.nV, m.nTri, m.nz)
    ub.check(L.ufe_b200_calc_secondary_velocities(ctx.h, None, None, C.byref(sec_s)))
    return state, vel, sec, nv.value, na.value


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("cfg")
    ap.add_argument("--resolution", type=float, default=None, help="mesh resolution in metres (default: maximum_resolution_uniform_config)")
    ap.add_argument("--backend", default=None, help="overrides choice_linear_algebra_backend_config")
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args(argv)
    values = cn.read_cfg(args.cfg)
    if args.backend:
        values["choice_linear_algebra_backend"] = args.backend
    setup = cn.run_setup(values)
    if not setup.use_b200:
        sys.exit("choice_linear_algebra_backend_config = 'PETSc': that is the reference's own CPU path, which is not part of this repository "
                 "(set the key to 'B200' or pass --backend B200)")
    if setup.approximation is None:
        sys.exit("choice_stress_balance_approximation_config = 'none': nothing to solve")
    w = workload_from_setup(setup, args.resolution)
    m = w.mesh
    ctx = ub.Context(device=args.device)          # fails loudly without a CUDA device: there is no CPU fallback
    try:
        ctx.mesh_create(m)
        ctx.calc_matrix_operators()
        state, vel, sec, n_visc, n_axb = solve_stress_balance(ctx, w, setup)
    finally:
        ctx.close()
    path = "UFEMISM/run_model(ANT)/run_ice_dynamics_model/run_ice_dynamics_model_pc/solve_stress_balance"
    print(f" mesh: nV = {m.nV}, nTri = {m.nTri}; n_visc_its = {n_visc}, n_Axb_its = {n_axb}")
    if setup.approximation in (ub.APPROX_SSA, ub.APPROX_DIVA):
        print(checksum_line("tau_dx_b", state["tau_dx_b"], path)); print(checksum_line("tau_dy_b", state["tau_dy_b"], path))
        print(checksum_line("u_b", state["u_vav_b"], path + "/relax_viscosity_iterations")); print(checksum_line("v_b", state["v_vav_b"], path + "/relax_viscosity_iterations"))
    for k in ("u_3D_b", "v_3D_b"):
        print(checksum_line("ice%" + k, vel[k], path))
    for k in ("u_surf", "v_surf", "u_vav", "v_vav", "uabs_surf", "uabs_vav"):
        print(checksum_line("ice%" + k, sec[k], path + "/calc_secondary_velocities"))


if __name__ == "__main__":
    main()
