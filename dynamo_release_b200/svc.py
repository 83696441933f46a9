"""SvcVectorField: drop-in for dynamo.vectorfield.scVectorField.SvcVectorField (reference :912-1164, with the
BaseVectorField / DifferentiableVectorField methods it inherits, :545-597, :787-909) on the B200 kernels.
"""
from __future__ import annotations

import numpy as np

from . import vf as _vf
from .sparsevfc import SparseVFC


def _update_dict(dict1, dict2):
    """tools/utils.py:125-137: only keys already present are overwritten."""
    dict1.update((k, dict2[k]) for k in dict1.keys() & dict2.keys())
    return dict1


def vecfld_from_adata(adata, basis="", vf_key="VecFld"):
    """reference vectorfield/utils.py:269-308: rebuild (vf_dict, func) from adata.uns['VecFld_<basis>']."""
    if basis is not None or len(basis) > 0:
        vf_key = "%s_%s" % (vf_key, basis)
    if vf_key not in adata.uns.keys():
        raise ValueError(
            f"Vector field function {vf_key} is not included in the adata object! "
            f"Try firstly running dyn.vf.VectorField(adata, basis='{basis}')"
        )
    vf_dict = adata.uns[vf_key]
    method = vf_dict.get("method", "sparsevfc")
    if method.lower() != "sparsevfc":
        raise ValueError(f"current only support two methods, SparseVFC and dynode: got {method}; only SparseVFC "
                         f"is served by the B200 backend")
    return vf_dict, _vf.RkhsField(vf_dict)


def _t_end(X, V):
    """tools/utils.py:3175-3189 (getTend): data extent over the 1st percentile of the non-zero speeds."""
    V_abs = np.abs(V)
    return np.max(X.max(0) - X.min(0)) / np.percentile(V_abs[V_abs > 0], 1)


def _t_seq(init_states, t_end, step_size=None):
    """tools/utils.py:3192-3209 (getTseq)."""
    if step_size is None:
        max_steps = int(max(7 / (init_states.shape[1] / 300), 4)) if init_states.shape[1] > 300 else 7
        return np.linspace(0, t_end, 10 ** (np.min([int(np.log10(t_end)), max_steps])))
    return np.arange(0, t_end + step_size, step_size)


class SvcVectorField:
    """Same constructor, attributes (`data`, `vf_dict`, `func`, `parameters`, `norm_dict`, `normalize`) and
    methods as the reference class."""

    def __init__(self, X=None, V=None, Grid=None, normalize=None, *args, **kwargs):
        self.data = {"X": X, "V": V, "Grid": Grid}
        self.vf_dict = kwargs.pop("vf_dict", {})
        self.func = kwargs.pop("func", None)
        self.fixed_points = kwargs.pop("fixed_points", None)
        if X is not None and V is not None:
            defaults = {
                "M": kwargs.pop("M", None) or max(min([50, len(X)]), int(0.05 * len(X)) + 1),  # :976
                "a": kwargs.pop("a", 5),
                "beta": kwargs.pop("beta", None),
                "ecr": kwargs.pop("ecr", 1e-5),
                "gamma": kwargs.pop("gamma", 0.9),
                "lambda_": kwargs.pop("lambda_", 3),
                "minP": kwargs.pop("minP", 1e-5),
                "MaxIter": kwargs.pop("MaxIter", 500),
                "theta": kwargs.pop("theta", 0.75),
                "div_cur_free_kernels": kwargs.pop("div_cur_free_kernels", False),
                "velocity_based_sampling": kwargs.pop("velocity_based_sampling", True),
                "sigma": kwargs.pop("sigma", 0.8),
                "eta": kwargs.pop("eta", 0.5),
                "seed": kwargs.pop("seed", 0),
            }
            self.parameters = {**kwargs, **defaults}
        self.norm_dict = {}
        self.normalize = normalize

    # ---- BaseVectorField (:577-597) ----
    def from_adata(self, adata, basis="", vf_key="VecFld"):
        vf_dict, func = vecfld_from_adata(adata, basis=basis, vf_key=vf_key)
        self.data["X"] = vf_dict["X"]
        self.data["V"] = vf_dict["Y"]  # use the raw velocity
        self.vf_dict = vf_dict
        self.func = func

    def get_X(self, idx=None):
        return self.data["X"] if idx is None else self.data["X"][idx]

    def get_V(self, idx=None):
        return self.data["V"] if idx is None else self.data["V"][idx]

    def get_data(self):
        return self.data["X"], self.data["V"]

    # ---- BaseVectorField.integrate (:721-784) ----
    def integrate(self, init_states, dims=None, scale=1, t_end=None, step_size=None, args=(),
                  integration_direction="forward", interpolation_num=250, average=True, sampling="arc_length",
                  verbose=False, disable=False):
        """Integrate initial states through the field on the device (K11).  Like the reference, `scale` only takes
        effect when `func` is still unset; `dims` must keep every dimension of the field (the kernel integrates the
        full RKHS field)."""
        from .integrate import ScaledRkhsField, integrate_vf_ivp

        init_states = np.asarray(init_states)
        if np.isscalar(dims):
            init_states = init_states[:, :dims]
        elif dims is not None:
            init_states = init_states[:, dims]
        if dims is not None and init_states.shape[1] != np.shape(self.vf_dict["X_ctrl"])[1]:
            raise NotImplementedError("integrate(dims=...) over a subset of the field's dimensions is not served by "
                                      "the B200 backend")
        if self.func is None:
            self.func = ScaledRkhsField(self.vf_dict, scale)
        if t_end is None:
            t_end = _t_end(self.get_X(), self.get_V())
        return integrate_vf_ivp(init_states, t=_t_seq(init_states, t_end, step_size),
                                integration_direction=integration_direction, f=self.func, args=args,
                                interpolation_num=interpolation_num, average=average, sampling=sampling,
                                verbose=verbose, disable=disable)

    # ---- train (:997-1046) ----
    def train(self, **kwargs):
        if self.normalize:
            # the reference's normalize=True path returns a dict without X_ctrl / C (denorm, :138-182) and then
            # fails inside vector_field_function; nothing well-defined to reproduce.
            raise NotImplementedError("normalize=True is not supported by the B200 backend (default is False)")
        verbose = kwargs.pop("verbose", 0)
        lstsq_method = kwargs.pop("lstsq_method", "drouin")
        known = ("M", "a", "beta", "ecr", "gamma", "lambda_", "minP", "MaxIter", "theta", "div_cur_free_kernels",
                 "velocity_based_sampling", "sigma", "eta", "seed", "sharded", "gather")
        params = {k: v for k, v in self.parameters.items() if k in known}
        VecFld = SparseVFC(self.data["X"], self.data["V"], self.data["Grid"], **params, verbose=verbose,
                           lstsq_method=lstsq_method)
        self.parameters = _update_dict(self.parameters, VecFld)
        self.vf_dict = VecFld
        self.func = _vf.RkhsField(VecFld)
        # :1043 evaluates f on ALL rows and overwrites the EM's V.  When every row is valid that is the same
        # product U C the loop just left in V, so it is only recomputed when rows were dropped.
        # (VecFld["X"] is the copy of data["X"]; in a sharded fit with gather it holds the rows of ALL ranks on rank 0)
        if len(VecFld["valid_ind"]) != len(VecFld["X"]):
            self.vf_dict["V"] = self.func(VecFld["X"])
        self.vf_dict["normalize"] = self.normalize
        return self.vf_dict

    # ---- DifferentiableVectorField (:787-909) ----
    def get_Jacobian(self, method="analytical", input_vector_convention="row", **kwargs):
        if method in ("analytical", "parallel"):
            kwargs.pop("cores", None)
            return _vf.RkhsJacobian(self.vf_dict, **kwargs)
        if method == "numerical":
            raise NotImplementedError("numerical Jacobians (numdifftools) are outside the B200 hot path; "
                                      "use method='analytical'")
        raise NotImplementedError(
            f"The method {method} is not implemented. Currently only "
            f"supports 'analytical', 'numerical', and 'parallel'."
        )

    def compute_divergence(self, X=None, method="analytical", **kwargs):
        X = self.data["X"] if X is None else X
        return _vf.compute_divergence(self.get_Jacobian(method=method), X, **kwargs)

    def compute_acceleration(self, X=None, method="analytical", **kwargs):
        X = self.data["X"] if X is None else X
        return _vf.compute_acceleration(self.func, self.get_Jacobian(method=method), X, **kwargs)

    def compute_curvature(self, X=None, method="analytical", formula=2, **kwargs):
        X = self.data["X"] if X is None else X
        return _vf.compute_curvature(self.func, self.get_Jacobian(method=method), X, formula=formula, **kwargs)

    def compute_curl(self, X=None, method="analytical", dim1=0, dim2=1, dim3=2, **kwargs):
        """reference scVectorField.py:806-838 (the reference slices X to the chosen 2 / 3 dimensions before it calls
        the Jacobian, so like there this only works for 2- or 3-dimensional fields)."""
        X = self.data["X"] if X is None else X
        X = X[:, [dim1, dim2]] if (dim3 is None or X.shape[1] < 3) else X[:, [dim1, dim2, dim3]]
        return _vf.compute_curl(self.get_Jacobian(method=method, **kwargs), X, **kwargs)

    def compute_torsion(self, X=None, method="analytical", **kwargs):
        """reference scVectorField.py:877-892."""
        X = self.data["X"] if X is None else X
        return _vf.compute_torsion(self.func, self.get_Jacobian(method=method), X, **kwargs)

    def compute_sensitivity(self, X=None, method="analytical", **kwargs):
        """reference scVectorField.py:894-909."""
        X = self.data["X"] if X is None else X
        return _vf.compute_sensitivity(self.get_Jacobian(method=method), X, **kwargs)

    def get_Hessian(self, method="analytical", **kwargs):
        """reference scVectorField.py:1090-1112."""
        if method == "analytical":
            return lambda x: _vf.Hessian_rkhs_gaussian(x, self.vf_dict, **kwargs)
        if method == "numerical":
            raise Exception("numerical Hessian for vector field is not defined.")
        raise NotImplementedError(f"The method {method} is not implemented. Currently only supports 'analytical'.")

    def get_Laplacian(self, method="analytical", **kwargs):
        """reference scVectorField.py:1114-1135."""
        if method == "analytical":
            return lambda x: _vf.Laplacian(H=x)
        if method == "numerical":
            raise Exception("Numerical Laplacian for vector field is not defined.")
        raise NotImplementedError(f"The method {method} is not implemented. Currently only supports 'analytical'.")
