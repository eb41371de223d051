from .constants import (NEBULA, NEBULA_ENABLE_NEBULA_LOAD, NEBULA_ENABLE_NEBULA_LOAD_DEFAULT, NEBULA_ENABLED,
                        NEBULA_ENABLED_DEFAULT, NEBULA_LOAD_PATH, NEBULA_LOAD_PATH_DEFAULT,
                        NEBULA_NUM_OF_VERSION_IN_RETENTION, NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT,
                        NEBULA_PERSISTENT_STORAGE_PATH, NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT,
                        NEBULA_PERSISTENT_TIME_INTERVAL, NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT)


class DeepSpeedNebulaConfig:
    """``nebula`` section of the ds_config (reference ``nebula/config.py``)."""

    def __init__(self, param_dict):
        d = param_dict.get(NEBULA, {}) if NEBULA in param_dict else {}
        self.enabled = d.get(NEBULA_ENABLED, NEBULA_ENABLED_DEFAULT)
        self.load_path = d.get(NEBULA_LOAD_PATH, NEBULA_LOAD_PATH_DEFAULT)
        self.enable_nebula_load = d.get(NEBULA_ENABLE_NEBULA_LOAD, NEBULA_ENABLE_NEBULA_LOAD_DEFAULT)
        self.persistent_storage_path = d.get(NEBULA_PERSISTENT_STORAGE_PATH, NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT)
        self.persistent_time_interval = d.get(NEBULA_PERSISTENT_TIME_INTERVAL, NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT)
        self.num_of_version_in_retention = d.get(NEBULA_NUM_OF_VERSION_IN_RETENTION,
                                                 NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT)

    def repr(self):
        return self.__dict__
