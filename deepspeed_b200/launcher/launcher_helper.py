"""Environment translation for MPI-style launchers (reference ``launcher/launcher_helper.py``): the rank variables an
``mpirun`` / ``srun`` job exports become ``RANK`` / ``LOCAL_RANK``."""


def _agree(env, names, what):
    found = {env[n] for n in names or () if n in env}
    if not found:
        raise EnvironmentError(f"{what} number is not in current env!")
    if len(found) > 1:
        raise EnvironmentError(f"{what} number doesn't match!")
    return found.pop()


def env_mapping(env, rank_name_list=None, local_rank_name_list=None):
    env["RANK"] = _agree(env, rank_name_list, "rank")
    env["LOCAL_RANK"] = _agree(env, local_rank_name_list, "local_rank")
    return env
