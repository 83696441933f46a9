"""install(): rebind dynamo's hot-path names to the B200 implementations.

dynamo looks these functions up by module attribute, and several modules did `from .utils import ...` at import
time, so every module that holds its own reference is patched (scVectorField.py:38-61, vector_calculus.py:37-48,
VectorField.py).  `uninstall()` restores the originals.  Nothing is patched unless `dynamo` is importable."""
from __future__ import annotations

import importlib

from . import integrate, lap, rank, svc, vf
from .sparsevfc import SparseVFC

_NAMES = {
    "con_K": vf.con_K,
    "vector_field_function": vf.vector_field_function,
    "Jacobian_rkhs_gaussian": vf.Jacobian_rkhs_gaussian,
    "Jacobian_rkhs_gaussian_parallel": vf.Jacobian_rkhs_gaussian_parallel,
    "compute_divergence": vf.compute_divergence,
    "compute_acceleration": vf.compute_acceleration,
    "compute_curvature": vf.compute_curvature,
    "elementwise_jacobian_transformation": vf.elementwise_jacobian_transformation,
    "subset_jacobian_transformation": vf.subset_jacobian_transformation,
    "transform_jacobian": vf.transform_jacobian,
    "vector_transformation": vf.vector_transformation,
    "Hessian_rkhs_gaussian": vf.Hessian_rkhs_gaussian,
    "Laplacian": vf.Laplacian,
    "hessian_transformation": vf.hessian_transformation,
    "elementwise_hessian_transformation": vf.elementwise_hessian_transformation,
    "compute_curl": vf.compute_curl,
    "compute_torsion": vf.compute_torsion,
    "compute_sensitivity": vf.compute_sensitivity,
    "vecfld_from_adata": svc.vecfld_from_adata,
    "integrate_vf_ivp": integrate.integrate_vf_ivp,
    "_fate": integrate._fate,
    "rank_jacobian_genes": rank.rank_jacobian_genes,
    "action": lap.action,
    "action_aux": lap.action_aux,
    "action_grad": lap.action_grad,
    "action_grad_aux": lap.action_grad_aux,
    "least_action_path": lap.least_action_path,
    "SparseVFC": SparseVFC,
    "SvcVectorField": svc.SvcVectorField,
}
_MODULES = ("dynamo.vectorfield.utils", "dynamo.vectorfield.scVectorField", "dynamo.vectorfield.vector_calculus",
            "dynamo.vectorfield.VectorField", "dynamo.vectorfield.rank_vf", "dynamo.vectorfield", "dynamo.vf", "dynamo.prediction.utils",
            "dynamo.prediction.fate", "dynamo.vectorfield.least_action_path", "dynamo.prediction.least_action_path")
# prediction/least_action_path.py has its OWN least_action_path (time-step optimisation around lap_T, :423-480) on top
# of the same action / action_grad: only the four functional names are rebound there, its driver stays
_SKIP = {("dynamo.prediction.least_action_path", "least_action_path")}
_saved = {}
_combined = {}


def _svc_class():
    """The class bound to the name SvcVectorField: this package's class with dynamo's own SvcVectorField as a second
    base, so that the methods outside the hot path (find_fixed_points, get_fixed_points, assign_fixed_points,
    construct_graph, plot_energy, evaluate: scVectorField.py:569-719, 1048-1164) keep working on the same
    `data` / `vf_dict` / `func` attributes; every hot-path method resolves to the B200 one first."""
    if "cls" in _combined:
        return _combined["cls"]
    try:
        ref = importlib.import_module("dynamo.vectorfield.scVectorField").SvcVectorField
    except Exception:
        return svc.SvcVectorField
    if isinstance(ref, type) and issubclass(ref, svc.SvcVectorField):   # already installed
        return ref
    _combined["cls"] = type("SvcVectorField", (svc.SvcVectorField, ref), {
        "__doc__": svc.SvcVectorField.__doc__, "__module__": svc.SvcVectorField.__module__})
    return _combined["cls"]


def install(modules=_MODULES):
    """Returns the list of (module, name) pairs that were rebound."""
    done = []
    names = dict(_NAMES)
    if not _saved:
        _combined.clear()
    names["SvcVectorField"] = _svc_class()
    for modname in modules:
        try:
            mod = importlib.import_module(modname)
        except Exception:
            continue
        for name, fn in names.items():
            if (modname, name) in _SKIP:
                continue
            if hasattr(mod, name) and getattr(mod, name) is not fn:
                _saved[(modname, name)] = getattr(mod, name)
                setattr(mod, name, fn)
                done.append((modname, name))
    return done


def uninstall():
    for (modname, name), orig in list(_saved.items()):
        setattr(importlib.import_module(modname), name, orig)
        del _saved[(modname, name)]
    _combined.clear()
