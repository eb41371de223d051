from .profiler import FlopsProfiler, get_model_profile, number_to_string, duration_to_string
