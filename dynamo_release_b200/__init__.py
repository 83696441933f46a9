"""dynamo_release_b200 -- B200-native (sm_100a) implementation of dynamo's SparseVFC vector-field hot path.

Drop-in names (same signatures as dynamo.vectorfield.scVectorField / dynamo.vectorfield.utils):
    SparseVFC, SvcVectorField, con_K, vector_field_function, Jacobian_rkhs_gaussian,
    Jacobian_rkhs_gaussian_parallel, compute_divergence, compute_acceleration, compute_curvature,
    elementwise_jacobian_transformation, subset_jacobian_transformation, transform_jacobian,
    vector_transformation, vecfld_from_adata, integrate_vf_ivp (dynamo.prediction.utils)
`install()` (patch.py) rebinds them inside an importable dynamo.  There is no CPU fallback.
"""
from .sparsevfc import SparseVFC, fit_em  # noqa: F401
from .svc import SvcVectorField, vecfld_from_adata  # noqa: F401
from .integrate import (  # noqa: F401
    DenseSolution, ScaledRkhsField, _fate, dense_eval_many, integrate_vf_ivp, solve_ivp_batch,
)
from .lap import action, action_grad, least_action_path  # noqa: F401
from .rank import rank_jacobian_genes  # noqa: F401
from .vf import (  # noqa: F401
    Hessian_rkhs_gaussian,
    Laplacian,
    compute_curl,
    compute_sensitivity,
    compute_torsion,
    elementwise_hessian_transformation,
    hessian_transformation,
    laplacian_rkhs,
    Jacobian_rkhs_gaussian,
    Jacobian_rkhs_gaussian_parallel,
    RkhsField,
    RkhsJacobian,
    average_jacobian_by_group,
    cell_angles,
    compute_acceleration,
    compute_curvature,
    compute_divergence,
    con_K,
    elementwise_jacobian_transformation,
    jacobian_det,
    jacobian_det_rkhs,
    speed,
    subset_jacobian_transformation,
    transform_jacobian,
    vector_field_function,
    vector_transformation,
)

__version__ = "0.1.0"
