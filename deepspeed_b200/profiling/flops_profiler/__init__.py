from .profiler import (FlopsProfiler, bytes_to_string, duration_to_string, flops_to_string, get_model_profile,  # noqa: F401
                       get_module_duration, get_module_flops, get_module_macs, macs_to_string, number_to_string,
                       params_to_string, wrapFunc)
