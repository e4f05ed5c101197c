"""`keypoint_moseq.initialize` of the reference, served by keypointmoseq_b200.initialize (ref: initialize.py:20-64)."""
from keypointmoseq_b200.initialize import *  # noqa: F401,F403
from keypointmoseq_b200.initialize import (initial_ar_params, initial_hdp_transitions, initial_heading,  # noqa: F401
                                           initial_latents, initial_location, initial_transitions)
