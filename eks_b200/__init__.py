"""eks_b200 -- B200-native (sm_100a CUDA) implementation of the hot path of erialc-cal/eks:
ensemble reduction -> Kalman filter -> RTS smoother (+ Newton-recursive / block-tridiagonal
variants), behind the reference's own Python function names.

    eks_b200.ensemble_kalman         drop-in for eks/ensemble_kalman.py   (numpy in / numpy out)
    eks_b200.newton_eks              kalman_newton_recursive               (eks/newton_eks.py:115)
    eks_b200.device                  batched device-tensor API over the C ABI (include/eksb.h)

No CPU fallback exists anywhere in this package.
"""
__version__ = "0.1.0"
