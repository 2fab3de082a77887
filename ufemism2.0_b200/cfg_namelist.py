"""The reference's .cfg files driving this path: a reader for the `&CONFIG ... /` Fortran namelist UFEMISM is configured with, and the mapping of
the keys the hot path reads onto ufe_b200_config -- the Python twin of ufe_b200_fill_config (fortran/ufe_b200_mod.f90), so that the ctypes driver and the
tests are set up from the same text a UFEMISM run is.

Reference behaviour mirrored here (src/UPSY/basic/model_configuration/):
  * every key ends in `_config`; the struct member is the key without that suffix (model_configuration_type_and_namelist.f90, copy_config_variables_to_struct);
  * a key the model does not know stops the run (model_configuration_utilities.f90:36-43) -- here: a key of *this path* with a value it does not know
    raises, keys of other sub-models are carried along untouched;
  * the new keys of INTEGRATION.md: choice_linear_algebra_backend_config ('PETSc' | 'B200'), B200_KSP_type_config, B200_PC_type_config,
    B200_use_initial_guess_config, B200_stiffness_format_config.
"""
import re

import ufe_b200 as ub


class ConfigError(ValueError):
    pass


def _scalar(text):
    t = text.strip()
    if t[:1] in "'\"":
        q = t[0]
        end = t.find(q, 1)
        if end < 0:
            raise ConfigError(f"unterminated string: {text!r}")
        return t[1:end]
    t = t.split("!")[0].strip().rstrip(",")
    u = t.upper()
    # Fortran list-directed input reads a logical from its first letter after an optional period (the reference's own files hold `.FALSE`)
    if u[:2] == ".T" or u[:1] == "T":
        return True
    if u[:2] == ".F" or u[:1] == "F":
        return False
    try:
        return int(t)
    except ValueError:
        pass
    try:
        return float(u.replace("D", "E").replace("_DP", ""))
    except ValueError as e:
        raise ConfigError(f"cannot read the value {text!r}") from e


def _value(text):
    """one value, or a list for the array-valued keys (comma-separated; commas and `!` inside quotes belong to the string)"""
    parts, cur, quote = [], "", None
    for ch in text:
        if quote:
            cur += ch
            if ch == quote:
                quote = None
        elif ch in "'\"":
            quote = ch; cur += ch
        elif ch == "!":
            break
        elif ch == ",":
            parts.append(cur); cur = ""
        else:
            cur += ch
    if cur.strip():
        parts.append(cur)
    vals = [_scalar(p) for p in parts if p.strip()]
    if not vals:
        raise ConfigError(f"cannot read the value {text!r}")
    return vals[0] if len(vals) == 1 else vals


def parse_namelist(text):
    """{key without the _config suffix: value} of the `&CONFIG` group; comments (`!`) and blank lines are skipped, later assignments win (as in Fortran)."""
    out = {}
    inside = False
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line.startswith("!"):
            continue
        if line.upper().startswith("&CONFIG"):
            inside = True
            continue
        if line.startswith("/"):
            break
        if not inside:
            continue
        m = re.match(r"([A-Za-z0-9_+\-]+)\s*=\s*(.*)$", line)
        if not m:
            raise ConfigError(f"cannot read the namelist line {raw!r}")
        key = m.group(1)
        if not key.endswith("_config"):
            raise ConfigError(f"namelist key {key!r} does not end in _config")
        out[key[:-len("_config")]] = _value(m.group(2))
    return out


def read_cfg(path):
    with open(path) as f:
        return parse_namelist(f.read())


_BC_UV = {"infinite": ub.BC_INFINITE, "zero": ub.BC_ZERO, "periodic_ISMIP-HOM": ub.BC_PERIODIC_ISMIP_HOM, "infinite_SSA_icestream": ub.BC_INFINITE_SSA_ICESTREAM}
_BC_H = {"infinite": ub.BC_INFINITE, "zero": ub.BC_ZERO}
_SLIDING = {"no_sliding": ub.SLID_NO_SLIDING, "Weertman": ub.SLID_WEERTMAN, "Coulomb": ub.SLID_COULOMB, "Budd": ub.SLID_BUDD, "Tsai2015": ub.SLID_TSAI2015,
            "Schoof2005": ub.SLID_SCHOOF2005, "Zoet-Iverson": ub.SLID_ZOET_IVERSON}
_PC = {"jacobi": ub.PC_JACOBI, "none": ub.PC_NONE, "multigrid": ub.PC_MG}
_APPROX = {"SSA": ub.APPROX_SSA, "DIVA": ub.APPROX_DIVA, "SIA": ub.APPROX_SIA, "SIA/SSA": ub.APPROX_SIASSA}
_HYBRID = {"add": ub.HYBRID_ADD, "add_SIA_reduced": ub.HYBRID_ADD_SIA_REDUCED}
_SCALARS = ["visc_it_norm_dUV_tol", "visc_it_relax", "visc_eff_min", "vel_max", "stress_balance_PETSc_rtol", "stress_balance_PETSc_abstol",
            "Glens_flow_law_exponent", "Glens_flow_law_epsilon_sq_0", "slid_delta_v", "slid_Weertman_m", "slid_Budd_q_plastic", "slid_Budd_u_threshold",
            "slid_ZI_ut", "slid_ZI_p", "slid_beta_max", "subgrid_friction_exponent_on_B_grid", "Hi_min", "uniform_Glens_flow_factor",
            "refgeo_idealised_SSA_icestream_Hi", "refgeo_idealised_SSA_icestream_dhdx", "refgeo_idealised_SSA_icestream_L", "refgeo_idealised_SSA_icestream_m",
            "refgeo_idealised_ISMIP_HOM_L", "dHi_semiimplicit_fs", "dHi_PETSc_rtol", "dHi_PETSc_abstol", "dt_ice_min", "dt_ice_max"]
_SIDES = ["north", "east", "south", "west"]          # order of BC_u / BC_v / BC_H in ufe_b200_config


def _choice(table, value, what):
    if value not in table:
        raise ConfigError(f'unknown {what} "{value}"!')          # the reference's crash message
    return table[value]


class RunSetup:
    """What a .cfg selects for this path: the ufe_b200_config struct plus the choices that travel as call arguments."""

    def __init__(self, cfg, backend, approximation, hybrid_scheme, do_subgrid_friction_on_A_grid, values):
        self.cfg, self.backend, self.approximation, self.hybrid_scheme = cfg, backend, approximation, hybrid_scheme
        self.do_subgrid_friction_on_A_grid, self.values = do_subgrid_friction_on_A_grid, values

    @property
    def use_b200(self):
        return self.backend == "B200"


def run_setup(values):
    """values: parse_namelist output (keys not given keep the reference's namelist defaults)."""
    c = ub.default_config()
    for k in _SCALARS:
        if k in values:
            v = values[k]
            if isinstance(v, bool) or not isinstance(v, (int, float)):
                raise ConfigError(f"{k}_config must be a number, got {v!r}")
            setattr(c, k, float(v))
    if "visc_it_nit" in values:
        c.visc_it_nit = int(values["visc_it_nit"])
    for flag in ("do_include_SSADIVA_crossterms", "do_GL_subgrid_friction"):
        if flag in values:
            setattr(c, flag, 1 if values[flag] else 0)
    for k, side in enumerate(_SIDES):
        if f"BC_u_{side}" in values:
            c.BC_u[k] = _choice(_BC_UV, values[f"BC_u_{side}"], "choice_BC_u")
        if f"BC_v_{side}" in values:
            c.BC_v[k] = _choice(_BC_UV, values[f"BC_v_{side}"], "choice_BC_v")
        if f"BC_H_{side}" in values:
            c.BC_H[k] = _choice(_BC_H, values[f"BC_H_{side}"], "BC_H")
        if c.BC_v[k] == ub.BC_INFINITE_SSA_ICESTREAM:
            raise ConfigError('unknown choice_BC_v "infinite_SSA_icestream"!')      # offered for u only (solve_linearised_SSA_DIVA_infinite_slab.f90:583-641)
    law = values.get("choice_sliding_law")
    if law == "idealised":
        if values.get("choice_idealised_sliding_law") != "SSA_icestream":
            raise ConfigError(f'idealised sliding law "{values.get("choice_idealised_sliding_law")}" is not on this path')
        c.choice_sliding_law = ub.SLID_IDEALISED_SSA_ICESTREAM
    elif law is not None:
        c.choice_sliding_law = _choice(_SLIDING, law, "choice_sliding_law")
    if values.get("choice_ice_integration_method", "semi-implicit") != "none":      # 'none': the thickness does not evolve (calc_dHi_dt is not called)
        c.choice_ice_integration_method = _choice({"explicit": ub.INTEGRATION_EXPLICIT, "semi-implicit": ub.INTEGRATION_SEMI_IMPLICIT},
                                                  values.get("choice_ice_integration_method", "semi-implicit"), "choice_ice_integration_method")
    # the new keys
    backend = values.get("choice_linear_algebra_backend", "PETSc")
    if backend not in ("PETSc", "B200"):
        raise ConfigError(f'unknown choice_linear_algebra_backend "{backend}"!')
    c.ksp_type = int(values.get("B200_KSP_type", ub.KSP_BICGSTAB))
    pc = values.get("B200_PC_type", ub.PC_JACOBI)
    c.pc_type = _choice(_PC, pc, "B200_PC_type") if isinstance(pc, str) else int(pc)
    if c.ksp_type != ub.KSP_BICGSTAB or c.pc_type not in (ub.PC_JACOBI, ub.PC_NONE, ub.PC_MG):
        raise ConfigError(f"unknown B200_KSP_type / B200_PC_type {c.ksp_type} / {c.pc_type}")
    c.ksp_initial_guess_nonzero = 1 if values.get("B200_use_initial_guess", False) else 0
    c.stiffness_format = int(values.get("B200_stiffness_format", 0))
    c.write_diagnostics = 1 if values.get("do_write_checksum_log", True) else 0
    approx = values.get("choice_stress_balance_approximation", "DIVA")
    approximation = _choice(_APPROX, approx, "choice_stress_balance_approximation") if approx != "none" else None
    hybrid = _choice(_HYBRID, values.get("choice_hybrid_SIASSA_scheme", "add"), "choice_hybrid_SIASSA_scheme_config")
    return RunSetup(c, backend, approximation, hybrid, bool(values.get("do_subgrid_friction_on_A_grid", False)), values)


def run_setup_from_cfg(path):
    return run_setup(read_cfg(path))
