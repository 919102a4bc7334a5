"""`gradcheck` on an autograd Function class (reference: utils/gradcheck.py:34-57)."""
from torch.autograd.gradcheck import gradcheck as _gradcheck


def gradcheck(func, inputs, eps: float = 1e-6, atol: float = 1e-5, rtol: float = 1e-3,
              raise_exception: bool = True, check_sparse_nnz: bool = False, nondet_tol: float = 0.0,
              check_undefined_grad: bool = True, check_grad_dtypes: bool = False) -> bool:
    # `check_sparse_nnz` was removed from torch.autograd.gradcheck; accepted and ignored
    return _gradcheck(lambda *x: func.apply(*x), inputs, eps=eps, atol=atol, rtol=rtol,
                      raise_exception=raise_exception, nondet_tol=nondet_tol,
                      check_undefined_grad=check_undefined_grad, check_grad_dtypes=check_grad_dtypes)
