"""`LCFModel` of /root/reference/gp_reward-priors/optbnn/gp/models/model.py:202-513 behind the same constructor
(`LCFModel(p_covariance, function_vect, device, p_mean=None, name=None)`): a Gaussian process written as a linear
combination of functions, f(x) = Phi(x)^T w with w ~ N(p_mean, p_covariance), i.e. m(x) = Phi(x)^T p_mean and
K(x, x') = Phi(x)^T p_covariance Phi(x').  The feature map `function_vect` stays the caller's Python callable (the
reference takes any `(X, device) -> (n, d)` or `(X, aux_X, device) -> (n, d)`); everything downstream of Phi(X) - the
d x d Woodbury system, alpha = (K + eps I)^-1 (f - m), the log-density - is `mr_fprior_kernel` in libiqlb.so
(`csrc/mr_train.cu`), called from `MRTrainerF` (mr_training.py).  This class only holds the prior's constants on the
device in the form that kernel reads (fp64, as the reference computes them: model.py:290-305).
"""
import numpy as np
import torch


class LCFModel:
    def __init__(self, p_covariance, function_vect, device="cuda", p_mean=None, name=None):
        self.name, self.function_vect = name, function_vect
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("iqlpref_b200 has no CPU path: LCFModel holds its constants on a CUDA device")
        cov = torch.as_tensor(np.asarray(p_covariance) if not torch.is_tensor(p_covariance) else p_covariance)
        if cov.ndim == 1:  # model.py:236-241: a vector is the diagonal
            cov = torch.diag(cov)
        cov = cov.float()  # the reference keeps float32 tensors and widens per call (model.py:291)
        if p_mean is None:
            p_mean = torch.zeros(cov.shape[0])
        p_mean = torch.as_tensor(np.asarray(p_mean) if not torch.is_tensor(p_mean) else p_mean).float()
        if cov.ndim != 2 or cov.shape[0] != cov.shape[1] or p_mean.shape != (cov.shape[0],):
            raise ValueError("p_covariance must be (d,) or (d, d) and p_mean (d,)")
        self.n_concepts = int(cov.shape[0])
        self.p_mean = p_mean.to(self.device)
        self.p_covariance = cov.to(self.device)
        sigma = cov.double().cpu()
        # constants of the prior (the reference recomputes them every call: torch.linalg.inv, slogdet; model.py:300, 323)
        self._p_mean64 = p_mean.double().to(self.device).contiguous()
        self._sigma64 = sigma.to(self.device).contiguous()
        self._sigma_inv64 = torch.linalg.inv(sigma).to(self.device).contiguous()
        self._logdet_sigma = float(torch.linalg.slogdet(sigma)[1])

    def _feature_matrix(self, X, aux_X=None):  # model.py:255-260
        if aux_X is not None:
            return self.function_vect(X, aux_X, self.device)
        return self.function_vect(X, self.device)

    def features64(self, X, aux_X=None) -> torch.Tensor:
        """Phi(X) as the contiguous fp64 device matrix [n, d] the kernel reads."""
        phi = torch.as_tensor(self._feature_matrix(X, aux_X)).to(self.device, torch.float64).contiguous()
        if phi.ndim != 2 or phi.shape[1] != self.n_concepts:
            raise ValueError(f"function_vect returned {tuple(phi.shape)}, expected (n, {self.n_concepts})")
        return phi
