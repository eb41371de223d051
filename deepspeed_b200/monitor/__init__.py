from .monitor import MonitorMaster, TensorBoardMonitor, WandbMonitor, csvMonitor, CometMonitor


def _available(module, what):
    import importlib
    try:
        importlib.import_module(module)
        return True
    except ImportError:
        from ..utils.logging import logger
        logger.warning(f"{what} is not installed: that monitor is switched off")
        return False


def check_tb_availability():
    return _available("tensorboard", "tensorboard")


def check_wandb_availability():
    return _available("wandb", "wandb")


def check_comet_availability():
    return _available("comet_ml", "comet_ml")
