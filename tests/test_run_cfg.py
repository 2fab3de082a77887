"""run_cfg.py: the pieces of the .cfg-driven run that need no GPU -- the experiment a config file names, the refusal to run without the B200 switch, and
the checksum-log formatting that makes the output comparable with the reference's checksum_logfile.txt."""
import os

import numpy as np
import pytest

import cfg_namelist as cn
import run_cfg
import ufe_b200 as ub


def test_checksum_lines_are_formatted_like_the_reference_log():
    # values as they appear in the reference's logs (integrated_test_MISMIP_mod_small, integrated_test_SSA_icestream_small)
    for text in ("0.14803255706603E+02", "-.75954187690709E+03", "0.97174031812528E-02", "0.00000000000000E+00", "0.49437349185719E+07",
                 "-.10439152656247E-01", "0.10000000000000E+01"):
        x = float(text.replace("-.", "-0."))
        assert run_cfg.fortran_e(x) == text
    line = run_cfg.checksum_line("u_b", np.array([0.0, 0.25, -1.5]), "UFEMISM/solve_SSA")
    assert line == " u_b: sum = -.12500000000000E+01, sum(abs) = 0.17500000000000E+01, min = -.15000000000000E+01, max = 0.25000000000000E+00 [UFEMISM/solve_SSA]"


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree only exists in the build container")
def test_formatting_round_trips_every_number_of_a_reference_log():
    import re
    n = 0
    for l in open("/root/reference/automated_testing/UFEMISM/integrated_test_MISMIP_mod_small/reference/checksum_logfile.txt"):
        for tok in re.findall(r"= (-?[0-9]*\.[0-9]{14}E[+-][0-9]{2})", l):
            assert run_cfg.fortran_e(float(tok.replace("-.", "-0."))) == tok, tok
            n += 1
    assert n > 10000


GEOMETRIES = {
    "MISMIP_mod": ("xmin_ANT_config = -1000e3\nxmax_ANT_config = 1000e3\nrefgeo_idealised_MISMIP_mod_Hi_init_config = 100.0\n"
                   "choice_basal_hydrology_model_config = 'Martin2011'\nslid_ZI_phi_fric_uniform_config = 10.0\nmaximum_resolution_uniform_config = 100e3\n"
                   "choice_stress_balance_approximation_config = 'DIVA'\nchoice_mask_noice_config = 'MISMIP_mod'\n"),
    "SSA_icestream": ("xmin_ANT_config = -400e3\nxmax_ANT_config = 400e3\nmaximum_resolution_uniform_config = 64e3\n"
                      "choice_stress_balance_approximation_config = 'SSA'\nchoice_sliding_law_config = 'idealised'\n"
                      "choice_idealised_sliding_law_config = 'SSA_icestream'\nBC_u_west_config = 'infinite_SSA_icestream'\n"),
    "ISMIP-HOM_A": ("xmin_ANT_config = -160000.0\nxmax_ANT_config = 160000.0\nrefgeo_idealised_ISMIP_HOM_L_config = 160000.0\n"
                    "maximum_resolution_uniform_config = 16000.0\nchoice_sliding_law_config = 'no_sliding'\nBC_u_west_config = 'periodic_ISMIP-HOM'\n"),
    "Halfar": ("xmin_ANT_config = -1000e3\nxmax_ANT_config = 1000e3\nmaximum_resolution_uniform_config = 100e3\n"
               "choice_stress_balance_approximation_config = 'SIA/SSA'\nchoice_hybrid_SIASSA_scheme_config = 'add_SIA_reduced'\n"),
}


@pytest.mark.parametrize("geo", sorted(GEOMETRIES))
def test_workload_of_the_experiment_a_config_names(geo):
    text = "&CONFIG\nchoice_refgeo_init_idealised_config = '%s'\n%schoice_linear_algebra_backend_config = 'B200'\n/\n" % (geo, GEOMETRIES[geo])
    setup = cn.run_setup(cn.parse_namelist(text))
    assert setup.use_b200
    w = run_cfg.workload_from_setup(setup)
    m, f = w.mesh, w.f
    assert w.cfg is setup.cfg and w.approximation == setup.approximation
    assert abs(m.xmax - float(setup.values["xmax_ANT"])) < 1e-6 and m.nV > 100 and np.isfinite(f["Hi"]).all() and f["Hi"].max() > 0
    assert (f["mask_noice"].sum() > 0) == (geo == "MISMIP_mod")
    if geo == "MISMIP_mod":
        assert not f["Hi"][f["mask_noice"] == 1].any()                  # the initial slab ends where the mask begins (r = 900 km)
        assert set(np.unique(f["Hi"])) == {0.0, 100.0} and f["effective_pressure"].max() < 0.2 * wl_overburden(f)   # Martin2011 pore water almost floats the ice
    if geo == "ISMIP-HOM_A":
        assert abs(f["Hi"].mean() - 1000.0) < 20.0 and list(w.cfg.BC_u) == [0, 0, 0, 2]
    if geo == "SSA_icestream":
        assert (f["Hi"] == 2000.0).all() and w.cfg.choice_sliding_law == ub.SLID_IDEALISED_SSA_ICESTREAM
    if geo == "Halfar":
        assert abs(f["Hi"].max() - 3000.0) < 100.0 and setup.approximation == ub.APPROX_SIASSA and setup.hybrid_scheme == ub.HYBRID_ADD_SIA_REDUCED


def wl_overburden(f):
    return 917.0 * 9.81 * f["Hi"].max()


def test_without_the_switch_nothing_runs(tmp_path):
    p = tmp_path / "c.cfg"
    p.write_text("&CONFIG\nchoice_refgeo_init_idealised_config = 'Halfar'\n/\n")
    with pytest.raises(SystemExit, match="PETSc"):
        run_cfg.main([str(p)])
    p.write_text("&CONFIG\nchoice_refgeo_init_idealised_config = 'Bueler'\nchoice_linear_algebra_backend_config = 'B200'\n/\n")
    with pytest.raises(cn.ConfigError, match="Bueler"):
        run_cfg.main([str(p)])
