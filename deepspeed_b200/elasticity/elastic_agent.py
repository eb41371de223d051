"""``DSElasticAgent``: torch-elastic local agent that exports the DeepSpeed environment to its workers and restarts
the worker group when membership shrinks or grows (reference ``elasticity/elastic_agent.py:29``).

Behaviour kept: the agent's extra environment (``ds_env``) is merged into every worker's env; the rendezvous master
address defaults to this host's first IP; on a failed/unhealthy group, or when the rendezvous reports fewer
participants than at start, the group is restarted while restarts remain; when new nodes are waiting, the group is
restarted to admit them (scale-up). The batch size for the new world size comes from ``compute_elastic_config``."""
import os
import socket
import time
from contextlib import closing

from torch.distributed.elastic.agent.server.api import RunResult, WorkerState
from torch.distributed.elastic.agent.server.local_elastic_agent import LocalElasticAgent
from torch.distributed.elastic.metrics import put_metric
from torch.distributed.elastic.utils.logging import get_logger

log = get_logger(__name__)


class DSElasticAgent(LocalElasticAgent):

    def __init__(self, spec, env, start_method="spawn", exit_barrier_timeout: float = 300, log_dir=None):
        try:
            super().__init__(spec, start_method=start_method, exit_barrier_timeout=exit_barrier_timeout,
                             log_dir=log_dir)
        except TypeError:  # newer torch: logs are configured through a LogsSpecs object
            from torch.distributed.elastic.multiprocessing import DefaultLogsSpecs
            super().__init__(spec, logs_specs=DefaultLogsSpecs(log_dir=log_dir), start_method=start_method,
                             exit_barrier_timeout=exit_barrier_timeout)
        self.ds_env = dict(env)

    @staticmethod
    def _set_master_addr_port(store, master_addr, master_port, local_addr=None):
        if master_port is None:
            with closing(socket.socket(socket.AF_INET, socket.SOCK_STREAM)) as s:
                s.bind(("", 0))
                master_port = s.getsockname()[1]
        if master_addr is None:
            try:
                master_addr = socket.gethostbyname(socket.gethostname())
            except OSError:
                master_addr = "127.0.0.1"
        store.set("MASTER_ADDR", master_addr.encode("UTF-8"))
        store.set("MASTER_PORT", str(master_port).encode("UTF-8"))

    def _start_workers(self, worker_group):
        # export the DeepSpeed launcher environment to the workers that the base class spawns
        saved = {k: os.environ.get(k) for k in self.ds_env}
        os.environ.update({k: str(v) for k, v in self.ds_env.items()})
        try:
            return super()._start_workers(worker_group)
        finally:
            for k, v in saved.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v

    def _participants(self, rdzv_handler):
        state = getattr(getattr(rdzv_handler, "_state_holder", None), "state", None)
        return len(state.participants) if state is not None else None

    def _invoke_run(self, role: str = "default") -> RunResult:
        spec = self._worker_group.spec
        role = spec.role
        log.info(f"[{role}] starting workers for entrypoint: {spec.get_entrypoint_name()}")
        self._initialize_workers(self._worker_group)
        rdzv = spec.rdzv_handler
        at_start = self._participants(rdzv)
        while True:
            assert self._worker_group.state != WorkerState.INIT
            time.sleep(spec.monitor_interval)
            run_result = self._monitor_workers(self._worker_group)
            state = run_result.state
            self._worker_group.state = state
            put_metric(f"workers.{role}.remaining_restarts", self._remaining_restarts)
            put_metric(f"workers.{role}.{state.name.lower()}", 1)
            now = self._participants(rdzv)
            shrunk = at_start is not None and now is not None and now < at_start
            if state == WorkerState.SUCCEEDED:
                log.info(f"[{role}] worker group successfully finished.")
                self._exit_barrier()
                return run_result
            if state in (WorkerState.UNHEALTHY, WorkerState.FAILED) or shrunk:
                if self._remaining_restarts > 0:
                    log.info(f"[{role}] worker group {state.name} (participants {at_start} -> {now}); "
                             f"{self._remaining_restarts}/{spec.max_restarts} attempts left; restarting")
                    self._remaining_restarts -= 1
                    self._restart_workers(self._worker_group)
                    at_start = self._participants(rdzv)
                else:
                    self._stop_workers(self._worker_group)
                    self._worker_group.state = WorkerState.FAILED
                    self._exit_barrier()
                    return run_result
            elif state == WorkerState.HEALTHY:
                if rdzv.num_nodes_waiting() > 0:
                    log.info(f"[{role}] new nodes are waiting; restarting the worker group to admit them")
                    self._restart_workers(self._worker_group)
                    at_start = self._participants(rdzv)
            else:
                raise Exception(f"[{role}] Worker group in {state.name} state")
