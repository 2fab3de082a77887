"""The .cfg-driven run setup: UFEMISM's `&CONFIG` namelist read into ufe_b200_config (cfg_namelist.py, the twin of ufe_b200_fill_config in the Fortran
shim).  With the reference tree present (build container) the reference's own config files must give exactly the configurations the workloads use."""
import os

import pytest

import cfg_namelist as cn
import ufe_b200 as ub
import workloads as wl

REF = "/root/reference"

ICESTREAM = """
&CONFIG
    ! the keys of automated_testing/integrated_tests/idealised/SSA_icestream/config_01_32km.cfg that this path reads
    choice_stress_balance_approximation_config   = 'SSA'                            ! Choice of stress balance approximation
    visc_it_norm_dUV_tol_config                  = 5E-8
    visc_it_nit_config                           = 5000
    visc_it_relax_config                         = 0.3
    visc_eff_min_config                          = 1E4
    vel_max_config                               = 5000.0
    stress_balance_PETSc_rtol_config             = 1E-7
    stress_balance_PETSc_abstol_config           = 1E-5
    BC_u_west_config                             = 'infinite_SSA_icestream'         ! Boundary conditions
    BC_u_east_config                             = 'infinite_SSA_icestream'
    BC_u_south_config                            = 'zero'
    BC_u_north_config                            = 'zero'
    BC_v_west_config                             = 'zero'
    BC_v_east_config                             = 'zero'
    BC_v_south_config                            = 'zero'
    BC_v_north_config                            = 'zero'
    choice_sliding_law_config                    = 'idealised'
    choice_idealised_sliding_law_config          = 'SSA_icestream'
    Glens_flow_law_epsilon_sq_0_config           = 1E-10
    uniform_Glens_flow_factor_config             = 1E-18
    do_GL_subgrid_friction_config                = .TRUE.
    some_other_submodel_key_config               = 'carried along'
    choice_linear_algebra_backend_config         = 'B200'
    B200_PC_type_config                          = 'multigrid'
    B200_use_initial_guess_config                = .true.
/
"""


def same(a, b, skip=()):
    for name, _ in ub.Config._fields_:
        if name in skip or name == "reserved":
            continue
        x, y = getattr(a, name), getattr(b, name)
        if hasattr(x, "__len__"):
            assert list(x) == list(y), name
        else:
            assert x == y, (name, x, y)


def test_namelist_text_gives_the_workload_configuration():
    s = cn.run_setup(cn.parse_namelist(ICESTREAM))
    assert s.use_b200 and s.approximation == ub.APPROX_SSA and s.values["some_other_submodel_key"] == "carried along"
    assert s.cfg.pc_type == ub.PC_MG and s.cfg.ksp_initial_guess_nonzero == 1 and s.cfg.stiffness_format == 0
    same(s.cfg, wl.ssa_icestream(100e3).cfg, skip=("pc_type", "ksp_initial_guess_nonzero"))
    # defaults: the PETSc path stays selected, Jacobi, zero initial guess
    d = cn.run_setup({})
    assert not d.use_b200 and d.cfg.pc_type == ub.PC_JACOBI and d.cfg.ksp_initial_guess_nonzero == 0 and d.approximation == ub.APPROX_DIVA
    same(d.cfg, ub.default_config())


@pytest.mark.parametrize("line,msg", [("BC_u_west_config = 'periodic'", "unknown choice_BC_u"), ("BC_v_east_config = 'infinite_SSA_icestream'", "unknown choice_BC_v"),
                                      ("choice_sliding_law_config = 'Teflon'", "unknown choice_sliding_law"),
                                      ("choice_linear_algebra_backend_config = 'cuSPARSE'", "unknown choice_linear_algebra_backend"),
                                      ("B200_PC_type_config = 'ilu'", "unknown B200_PC_type"), ("visc_it_relax_config = 'a lot'", "must be a number"),
                                      ("choice_hybrid_SIASSA_scheme_config = 'multiply'", "unknown choice_hybrid_SIASSA_scheme_config"),
                                      ("visc_it_relax = 0.3", "does not end in _config")])
def test_unknown_choices_stop_the_run(line, msg):
    with pytest.raises(cn.ConfigError, match=msg):
        cn.run_setup(cn.parse_namelist("&CONFIG\n" + line + "\n/\n"))


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree only exists in the build container")
def test_reference_config_files_give_the_workload_configurations():
    # BASELINE.json configs[0]
    s = cn.run_setup_from_cfg(os.path.join(REF, "config-files", "config_MISMIP_8km_spinup_for_scaling.cfg"))
    assert s.approximation == ub.APPROX_DIVA and not s.use_b200                      # the reference's file knows nothing of the new switch
    unused = ("refgeo_idealised_SSA_icestream_Hi", "refgeo_idealised_SSA_icestream_dhdx", "refgeo_idealised_SSA_icestream_L", "refgeo_idealised_SSA_icestream_m")
    same(s.cfg, wl.scaling_config(), skip=unused)                                    # placeholders of an experiment this file does not run
    no_thickness = ("dHi_PETSc_rtol", "dHi_PETSc_abstol", "dt_ice_min", "BC_H")         # velocity-only experiments (choice_ice_integration_method = 'none')
    t = "automated_testing/integrated_tests/idealised/SSA_icestream/config_01_32km.cfg"
    s = cn.run_setup_from_cfg(os.path.join(REF, t))
    assert s.approximation == ub.APPROX_SSA
    same(s.cfg, wl.ssa_icestream(100e3).cfg, skip=no_thickness)
    t = "automated_testing/UFEMISM/integrated_test_ISMIP_HOM_small/all_ISMIP_HOM_config_files/config_ISMIP_HOM_A_160_DIVA.cfg"
    s = cn.run_setup_from_cfg(os.path.join(REF, t))
    assert s.approximation == ub.APPROX_DIVA and s.hybrid_scheme == ub.HYBRID_ADD
    same(s.cfg, wl.ismip_hom_a(h=40e3).cfg, skip=no_thickness + unused)
    t = "automated_testing/UFEMISM/integrated_test_MISMIP_mod_small/config.cfg"
    s = cn.run_setup_from_cfg(os.path.join(REF, t))
    from test_reference_logs import mismip_mod_small_workload
    same(s.cfg, mismip_mod_small_workload().cfg, skip=unused + ("dt_ice_min",))
    s = cn.run_setup_from_cfg(os.path.join(REF, "automated_testing/UFEMISM/integrated_test_SSA_icestream_small/config.cfg"))
    same(s.cfg, wl.ssa_icestream(100e3).cfg, skip=no_thickness)
