from .constants import *  # noqa: F401,F403

AIO_DEFAULT_DICT = {
    AIO_BLOCK_SIZE: AIO_BLOCK_SIZE_DEFAULT,
    AIO_QUEUE_DEPTH: AIO_QUEUE_DEPTH_DEFAULT,
    AIO_INTRA_OP_PARALLELISM: AIO_INTRA_OP_PARALLELISM_DEFAULT,
    AIO_SINGLE_SUBMIT: AIO_SINGLE_SUBMIT_DEFAULT,
    AIO_OVERLAP_EVENTS: AIO_OVERLAP_EVENTS_DEFAULT,
    AIO_USE_GDS: AIO_USE_GDS_DEFAULT,
}


def get_aio_config(param_dict):
    """The ``aio`` section merged over its defaults (reference ``aio_config.py:20``)."""
    given = param_dict.get(AIO) if isinstance(param_dict, dict) else None
    out = dict(AIO_DEFAULT_DICT)
    if given:
        out.update({k: v for k, v in given.items() if k in out})
    return out
