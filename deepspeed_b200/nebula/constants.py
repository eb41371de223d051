"""Keys of the ``nebula`` ds_config section (reference ``nebula/constants.py``). Nebula is Azure's asynchronous
checkpoint service; the engine needs the external ``torch_nebula`` package."""
NEBULA = "nebula"
NEBULA_ENABLED = "enabled"
NEBULA_ENABLED_DEFAULT = False
NEBULA_ENABLE_NEBULA_LOAD = "enable_nebula_load"
NEBULA_ENABLE_NEBULA_LOAD_DEFAULT = True
NEBULA_LOAD_PATH = "nebula_load_path"
NEBULA_LOAD_PATH_DEFAULT = None
NEBULA_PERSISTENT_STORAGE_PATH = "persistent_storage_path"
NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT = None
NEBULA_PERSISTENT_TIME_INTERVAL = "persistent_time_interval"
NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT = 100
NEBULA_NUM_OF_VERSION_IN_RETENTION = "num_of_version_in_retention"
NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT = 2
NEBULA_EXPORT_ENVS = ["DLTS_JOB_ID", "DLTS_NUM_WORKER", "NEBULA_PERSISTENT_STORAGE_PATH",
                      "NEBULA_PERSISTENT_TIME_INTERVAL", "AML_RUN_ID", "AZUREML_RUN_TOKEN",
                      "AZUREML_WORKSPACE_SCOPE", "AZUREML_EXPERIMENT_SCOPE", "AZUREML_RUN_HISTORY_SERVICE_ENDPOINT",
                      "AZUREML_RUN_ID", "NEBULA_MEMORY_BUFFER_SIZE", "AZUREML_PARAMETER_ITPJOB_NAME",
                      "FC_TASKROLE_NAME", "FC_TASK_INDEX", "MASTER_HOST", "LOCAL_HOST", "AZUREML_BLOB_ACCOUNT_NAME",
                      "AZUREML_BLOB_ACCOUNT_KEY"]
DLTS_POD_ENV_PATH = "/dlts-runtime/env/pod.env"
