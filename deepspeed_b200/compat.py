"""Helpers that hide differences between Python versions (reference ``deepspeed/compat.py``): annotation access moved to
``annotationlib`` in Python 3.14."""
from inspect import ismodule
from typing import Any, Dict, Mapping

try:
    import annotationlib  # Python >= 3.14
except ImportError:
    annotationlib = None


def get_annotations_from_namespace(namespace: Mapping[str, object]) -> Dict[str, Any]:
    """Annotations of a class body namespace (as seen by a metaclass)."""
    if annotationlib:
        fn = annotationlib.get_annotate_from_class_namespace(namespace)
        if fn is not None:
            return annotationlib.call_annotate_function(fn, annotationlib.Format.VALUE)
    return namespace.get("__annotations__", {})


def get_annotations(obj: Any) -> Dict[str, Any]:
    """Annotations of a module, class, function, or of the type of any other object."""
    if annotationlib:
        annotated = hasattr(obj, "__annotations__") or hasattr(obj, "__annotate__")
        if not isinstance(obj, type) and not ismodule(obj) and not callable(obj) and not annotated:
            obj = type(obj)
        return annotationlib.get_annotations(obj)
    return getattr(obj, "__annotations__", {})
