"""`keypoint_moseq.transitions` of the reference, served by keypointmoseq_b200.gibbs (ref: transitions.py:7-44)."""
from keypointmoseq_b200.gibbs import sample_hdp_transitions, sample_ms, sample_transitions  # noqa: F401
