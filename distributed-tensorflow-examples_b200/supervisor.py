"""Chief-driven lifecycle: ``tf.train.Supervisor`` + ``managed_session`` re-thought.

Reference use (``encoder:147-161,176``; same in gan/lstm): ``Supervisor(is_chief=(task_index
== 0), logdir=model_dir, init_op, summary_op=None, saver, global_step, save_model_secs=60)``,
``with sv.managed_session(server.target, config) as sess``, ``sv.should_stop()``, ``sv.stop()``.

Behaviour kept:
* chief: declare the variables on their ps tasks, restore the newest checkpoint in ``logdir``
  if there is one, otherwise run the initialisers; then start the checkpoint timer (first
  save right away, then every ``save_model_secs``) and, in sync mode, the aggregation runner;
* non-chief: declare, then poll ``ready_op`` (count of uninitialised variables) every
  ``recovery_wait_secs`` until the chief is done;
* ``should_stop()`` turns true after ``request_stop()`` or an exception inside the session;
* chief also writes ``graph.json`` (placement table, the analogue of ``graph.pbtxt``).
One addition: the chief writes a final checkpoint when the session closes.
"""
from __future__ import annotations

import contextlib
import json
import os
import threading
import time
from typing import Callable, Dict, Optional

import torch

from .checkpoint import Saver, latest_checkpoint


class Supervisor:
    def __init__(self, is_chief: bool, logdir: str, store, registry, model, saver: Optional[Saver] = None,
                 save_model_secs: float = 60, recovery_wait_secs: float = 30.0, seed: int = 0,
                 sync: bool = False, replicas_to_aggregate: int = 0, reinit_on_join: bool = False,
                 init_values: Optional[Callable[[], Dict[str, torch.Tensor]]] = None):
        self.is_chief, self.logdir = is_chief, logdir
        self.store, self.registry, self.model = store, registry, model
        self.saver = saver or Saver()
        self.save_model_secs, self.recovery_wait_secs = save_model_secs, recovery_wait_secs
        self.seed, self.sync = seed, sync
        self.replicas_to_aggregate = replicas_to_aggregate
        self.reinit_on_join = reinit_on_join
        self._init_values = init_values or (lambda: registry.initial_values(seed))
        self._stop = threading.Event()
        self._threads = []
        self.restored_from: Optional[str] = None
        self.last_saved: Optional[str] = None
        self._save_lock = threading.Lock()
        self._final_done = False

    # ---- stop protocol -------------------------------------------------------------
    def should_stop(self) -> bool:
        return self._stop.is_set()

    def request_stop(self, exc: Optional[BaseException] = None) -> None:
        self._stop.set()

    def stop(self) -> None:
        """Idempotent (the reference calls it twice: leaving managed_session and ``encoder:176``)."""
        self._stop.set()
        had_runner = any(t.name == "SyncReplicasQueueRunner" for t in self._threads)
        for t in self._threads:
            t.join(timeout=10.0)
        self._threads = []
        if had_runner and hasattr(self.store, "release_sync_waiters"):
            self.store.release_sync_waiters()      # the aggregation runner is gone: unblock waiting workers

    # ---- session bring-up ------------------------------------------------------------
    def prepare_or_wait_for_session(self) -> None:
        st = self.store
        st.declare_all()
        if self.sync:
            st.sync_setup(self.is_chief)
        if self.is_chief:
            latest = latest_checkpoint(self.logdir)
            if latest is not None:
                values = self._init_values()
                values.update(Saver.restore(latest))
                self.restored_from = latest
            else:
                values = self._init_values()
            st.initialize(values)
            self._write_graph()
            self._start_services()
        else:
            while st.uninitialized() > 0:
                if self._stop.wait(self.recovery_wait_secs):
                    return
        if self.reinit_on_join:
            # reference quirk (encoder:160, gan:181): every worker re-runs init on entry
            st.initialize(self._init_values())

    def _write_graph(self) -> None:
        os.makedirs(self.logdir, exist_ok=True)
        table = [{"name": n, "shape": list(s), "device": d} for n, s, d in self.registry.placement_table()]
        with open(os.path.join(self.logdir, "graph.json"), "w") as f:
            json.dump({"model": self.model.name, "variables": table}, f, indent=1)

    def _start_services(self) -> None:
        if self.save_model_secs and self.save_model_secs > 0:
            t = threading.Thread(target=self._checkpoint_loop, name="SVTimerCheckpointThread", daemon=True)
            t.start()
            self._threads.append(t)
        if self.sync:
            t = threading.Thread(target=self._aggregate_loop, name="SyncReplicasQueueRunner", daemon=True)
            t.start()
            self._threads.append(t)
        t = threading.Thread(target=self._step_counter_loop, name="SVStepCounterThread", daemon=True)
        t.start()
        self._threads.append(t)

    def _step_counter_loop(self, period: float = 5.0) -> None:
        """The chief's ``global_step/sec`` summary (TF's SVStepCounterThread): one JSON line per
        period appended to ``<logdir>/events.jsonl``."""
        from .control.client import PsError
        path = os.path.join(self.logdir, "events.jsonl")
        last_t, last_step = time.time(), None
        while not self._stop.wait(period):
            try:
                step = self.store.global_step()
            except (PsError, ConnectionError, OSError):
                return
            now = time.time()
            if last_step is not None and now > last_t:
                with open(path, "a") as f:
                    f.write(json.dumps({"wall_time": now, "global_step": step,
                                        "global_step/sec": (step - last_step) / (now - last_t)}) + "\n")
            last_t, last_step = now, step

    def save_checkpoint(self) -> Optional[str]:
        with self._save_lock:
            values = self.store.read_all()
            step = int(values[self.model.global_step.name])
            meta = {"model": self.model.name, "placement": [[n, list(s), d] for n, s, d in self.registry.placement_table()]}
            self.last_saved = self.saver.save(values, self.logdir, step, meta)
            return self.last_saved

    def _checkpoint_loop(self) -> None:
        while True:
            try:
                self.save_checkpoint()
            except Exception as e:  # keep training alive; report
                print("checkpoint failed: %r" % (e,), flush=True)
            if self._stop.wait(self.save_model_secs):
                return

    def _aggregate_loop(self) -> None:
        r = self.replicas_to_aggregate or self.store.num_workers
        from .control.client import PsError
        while not self._stop.is_set():
            try:
                self.store.chief_aggregate_once(r)
            except (PsError, ConnectionError):
                return

    @contextlib.contextmanager
    def managed_session(self):
        self.prepare_or_wait_for_session()
        try:
            yield self.store
        except BaseException:
            self.request_stop()
            raise
        finally:
            self.final_checkpoint()
            self.stop()

    def final_checkpoint(self) -> Optional[str]:
        """Chief only, once: stop the checkpoint timer and save the current state.  The worker
        loop calls it BEFORE it enqueues its done token, because the ps tasks exit as soon as the
        last token arrives and the variables go with them."""
        if not self.is_chief or not self.save_model_secs or self._final_done:
            return None
        self._final_done = True
        try:
            self._stop.set()
            for t in self._threads:
                if t.name == "SVTimerCheckpointThread":
                    t.join(timeout=10.0)
            return self.save_checkpoint()
        except Exception as e:
            print("final checkpoint failed: %r" % (e,), flush=True)
            return None
