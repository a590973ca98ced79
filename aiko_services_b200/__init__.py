"""aiko_services_b200 -- B200 (sm_100a) implementation of aiko_services' media hot path.

Only the per-frame image / audio PipelineElement path of the reference is implemented here
(SURVEY.md section 8); see DESIGN.md for the boundary and INTEGRATION.md for how the
reference binds to it.  The CUDA library is mandatory: there is no CPU fallback.
"""
__version__ = "0.1.0"
