"""`keypoint_moseq.distributions` of the reference, served by keypointmoseq_b200.distributions
(ref: distributions.py:5-79; the HMM sampler of :46-79 is `sample_hmm_stateseqs`, batched over sessions)."""
from keypointmoseq_b200.distributions import *  # noqa: F401,F403
from keypointmoseq_b200.distributions import (sample_gamma, sample_invwishart, sample_mn, sample_mniw,  # noqa: F401
                                              sample_scaled_inv_chi2)
from keypointmoseq_b200.gibbs import sample_hmm_stateseqs  # noqa: F401
