"""Command-line task driver: ``cxxnet <config> [name=value ...]``.

Behavioural parity with ``CXXNetLearnTask`` (reference ``src/cxxnet_main.cpp:16-471``): tasks
``train`` / ``finetune`` / ``pred`` / ``extract``; model files ``<model_dir>/%04d.model``;
``continue=1`` resumes from the latest consecutive model; ``model_in`` infers the round from
the file name; ``save_model`` period; progress on stdout, metrics on stderr as
``[round]\\ttrain-error:...\\ttest-error:...``; ``test_io`` runs the data pipeline only.
"""
from __future__ import annotations

import os
import re
import struct
import sys
import time
from typing import List, Optional

import numpy as np

from .config import ConfigPairs, parse_cli_overrides, parse_config_file
from .io.data import IIterator, create_iterator
from .trainer import NetTrainer, create_net, parse_devices


class LearnTask:
    def __init__(self):
        self.task = "train"
        self.net_type = 0
        self.reset_net_type = -1
        self.net_trainer: Optional[NetTrainer] = None
        self.itr_train: Optional[IIterator] = None
        self.itr_pred: Optional[IIterator] = None
        self.itr_evals: List[IIterator] = []
        self.eval_names: List[str] = []
        self.cfg: ConfigPairs = []
        self.test_io = 0
        self.print_step = 100
        self.num_round = 10
        self.max_round = 1 << 31
        self.continue_training = 0
        self.save_period = 1
        self.start_counter = 0
        self.silent = 0
        self.device = "cpu"
        self.name_model_in = "NULL"
        self.name_model_dir = "models"
        self.name_pred = "pred.txt"
        self.extract_node_name = ""
        self.output_format = 1
        self.rank = 0

    # ------------------------------------------------------------------ entry
    def run(self, argv: List[str]) -> int:
        if len(argv) < 1:
            print("Usage: <config> [name=value ...]")
            return 0
        for k, v in parse_config_file(argv[0]):
            self.set_param(k, v)
        for k, v in parse_cli_overrides(argv[1:]):
            self.set_param(k, v)
        kind, devs = parse_devices(self.device)
        if kind == "gpu" and len(devs) > 1 and "RANK" not in os.environ:
            from .parallel.launch import relaunch_per_device
            return relaunch_per_device(argv, devs)
        self._init_distributed()
        self.init()
        if self.net_trainer is not None and self.net_trainer.is_server:
            # update_on_server = 1: this rank owns weights and updaters (parallel/ps_server.py) until the workers finish
            stats = self.net_trainer.serve()
            if not self.silent:
                print("parameter server %d: %s" % (self.rank, stats))
            return 0
        if not self.silent and self.rank == 0:
            print("initializing end, start working")
        if self.task in ("train", "finetune"):
            self.task_train()
        if self.task == "pred":
            self.task_predict()
        if self.task in ("extract", "extract_feature"):
            self.task_extract_feature()
        if self.task == "pred_raw":
            # Documented by the reference (doc/tasks.md, example/kaggle_bowl/pred.conf) although its driver never
            # dispatches it (cxxnet_main.cpp:77-79): one text row of raw top-node outputs per instance.
            self.extract_node_name = self.extract_node_name or "top[-1]"
            self.output_format = 1
            self.task_extract_feature()
        if self.net_trainer is not None:
            self.net_trainer.finish()
        self.close()
        return 0

    def _init_distributed(self):
        if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
            import torch
            import torch.distributed as dist
            if not dist.is_initialized():
                kind, devs = parse_devices(self.device)
                local = int(os.environ.get("LOCAL_RANK", "0"))
                # a parameter-server run talks point to point through the host (its server ranks own no GPU)
                ps = any(k == "update_on_server" and v.strip() == "1" for k, v in self.cfg)
                if kind == "gpu" and not ps:
                    torch.cuda.set_device(devs[local] if local < len(devs) else local)
                dist.init_process_group("nccl" if kind == "gpu" and not ps else "gloo")
            self.rank = dist.get_rank()

    def set_param(self, name: str, val: str) -> None:
        if val == "default":
            return
        if name == "net_type": self.net_type = int(val)
        if name == "reset_net_type": self.reset_net_type = int(val)
        if name == "print_step": self.print_step = int(val)
        if name == "continue": self.continue_training = int(val)
        if name == "save_model": self.save_period = int(val)
        if name == "start_counter": self.start_counter = int(val)
        if name == "model_in": self.name_model_in = val
        if name == "model_dir": self.name_model_dir = val
        if name == "num_round": self.num_round = int(val)
        if name == "max_round": self.max_round = int(val)
        if name == "silent": self.silent = int(val)
        if name == "task": self.task = val
        if name == "dev": self.device = val
        if name == "test_io": self.test_io = int(val)
        if name == "extract_node_name": self.extract_node_name = val
        if name == "output_format":
            self.output_format = 1 if val == "txt" else 0
        self.cfg.append((name, val))

    # ------------------------------------------------------------------ init
    def init(self) -> None:
        if self.task == "train" and self.continue_training:
            if self.sync_latest_model() == 0:
                raise RuntimeError("Init: Cannot find models for continue training. "
                                   "Please specify it by model_in instead.")
            print("Init: Continue training from round %d" % self.start_counter)
            if not self.net_trainer.is_server:
                self.create_iterators()
            return
        self.continue_training = 0
        if self.name_model_in == "NULL":
            assert self.task == "train", "must specify model_in if not training"
            self.net_trainer = self.create_net()
            self.net_trainer.init_model()
        elif self.task == "finetune":
            self.copy_model()
        else:
            self.load_model()
        if not self.net_trainer.is_server:
            self.create_iterators()

    def _model_path(self, counter: int) -> str:
        return "%s/%04d.model" % (self.name_model_dir, counter)

    def sync_latest_model(self) -> int:
        s = self.start_counter
        last = None
        while os.path.exists(self._model_path(s)):
            last = self._model_path(s)
            s += 1
        if last is None:
            return 0
        with open(last, "rb") as f:
            self.net_type = struct.unpack("<i", f.read(4))[0]
            self.net_trainer = self.create_net()
            self.net_trainer.load_model(f)
        self.start_counter = s
        return 1

    def load_model(self) -> None:
        base = os.path.basename(self.name_model_in)
        m = re.match(r"(\d+)", base)
        if "/" in self.name_model_in and m:
            self.start_counter = int(m.group(1))
        else:
            print("WARNING: Cannot infer start_counter from model name. Specify it in config if needed")
        with open(self.name_model_in, "rb") as f:
            self.net_type = struct.unpack("<i", f.read(4))[0]
            self.net_trainer = self.create_net()
            self.net_trainer.load_model(f)
        self.start_counter += 1

    def save_model(self) -> None:
        name = self._model_path(self.start_counter)
        self.start_counter += 1
        if self.save_period == 0 or self.start_counter % self.save_period != 0:
            return
        if self.rank != 0:
            # replicas are identical; only rank 0 writes (but every rank syncs its arena view)
            self.net_trainer.net.sync_weights_from_device()
            return
        os.makedirs(self.name_model_dir, exist_ok=True)
        with open(name, "wb") as f:
            f.write(struct.pack("<i", self.net_type))
            self.net_trainer.save_model(f)

    def create_net(self) -> NetTrainer:
        if self.reset_net_type != -1:
            self.net_type = self.reset_net_type
        net = create_net(self.net_type)
        for k, v in self.cfg:
            net.set_param(k, v)
        return net

    def copy_model(self) -> None:
        with open(self.name_model_in, "rb") as f:
            self.net_type = struct.unpack("<i", f.read(4))[0]
            self.net_trainer = self.create_net()
            self.net_trainer.copy_model_from(f)

    def create_iterators(self) -> None:
        flag = 0
        evname = ""
        itcfg: ConfigPairs = []
        defcfg: ConfigPairs = []
        for name, val in self.cfg:
            if name == "data":
                flag = 1
                continue
            if name == "eval":
                evname = val
                flag = 2
                continue
            if name == "pred":
                flag = 3
                self.name_pred = val
                continue
            if name == "iter" and val == "end":
                assert flag != 0, "wrong configuration file"
                if flag == 1 and self.task in ("train", "finetune"):
                    assert self.itr_train is None, "can only have one data"
                    self.itr_train = create_iterator(itcfg)
                if flag == 2 and self.task in ("train", "finetune"):
                    self.itr_evals.append(create_iterator(itcfg))
                    self.eval_names.append(evname)
                if flag == 3 and self.task in ("pred", "pred_raw", "extract", "extract_feature"):
                    assert self.itr_pred is None, "can only have one data:test"
                    self.itr_pred = create_iterator(itcfg)
                flag = 0
                itcfg = []
            if flag == 0:
                defcfg.append((name, val))
            else:
                itcfg.append((name, val))
        world = int(os.environ.get("WORLD_SIZE", "1")) if "RANK" in os.environ else 1
        if self.net_trainer is not None and self.net_trainer.ps_roles is not None:
            world = self.net_trainer.ps_roles.num_workers        # parameter-server run: the workers share the data
        for it in [self.itr_train, self.itr_pred] + self.itr_evals:
            if it is not None:
                for k, v in defcfg:
                    it.set_param(k, v)
                if world > 1:
                    # every rank walks the whole data set but only uses its rows of each batch: let the batch stage
                    # skip the decoding of the others (io/iterators.py BatchAdaptIterator)
                    it.set_param("slice_world", str(world))
                    it.set_param("slice_rank", str(self.rank))
                it.init()

    def close(self) -> None:
        for it in [self.itr_train, self.itr_pred] + self.itr_evals:
            if it is not None:
                it.close()

    # ------------------------------------------------------------------ tasks
    def task_predict(self) -> None:
        assert self.itr_pred is not None, "must specify a predict iterator to generate predictions"
        print("start predicting...")
        with open(self.name_pred, "w") as fo:
            self.itr_pred.before_first()
            while self.itr_pred.next():
                batch = self.itr_pred.value()
                pred = self.net_trainer.predict(batch)
                assert batch.num_batch_padd < batch.batch_size, "num batch pad must be smaller"
                for j in range(pred.shape[0] - batch.num_batch_padd):
                    fo.write("%g\n" % pred[j])
        print("finished prediction, write into %s" % self.name_pred)

    def task_extract_feature(self) -> None:
        assert self.itr_pred is not None, "must specify a predict iterator to generate predictions"
        if not self.extract_node_name:
            raise RuntimeError("extract node name must be specified in task extract_feature.")
        print("start predicting...")
        nrow = 0
        dshape = (0, 0, 0)
        start = time.time()
        cnt = 0
        with open(self.name_pred, "wb") as fo:
            self.itr_pred.before_first()
            while self.itr_pred.next():
                batch = self.itr_pred.value()
                pred = self.net_trainer.extract_feature(batch, self.extract_node_name)
                assert batch.num_batch_padd < batch.batch_size, "num batch pad must be smaller"
                sz = pred.shape[0] - batch.num_batch_padd
                nrow += sz
                for j in range(sz):
                    d = pred[j].reshape(pred.shape[1], -1)
                    if self.output_format:
                        fo.write((" ".join("%g" % v for v in d.reshape(-1)) + " \n").encode())
                    else:
                        fo.write(np.ascontiguousarray(d, dtype="<f4").tobytes())
                if sz != 0:
                    dshape = tuple(pred.shape[1:])
                cnt += 1
                if cnt % self.print_step == 0 and not self.silent:
                    print("\rbatch:[%8d] %d sec elapsed" % (cnt, int(time.time() - start)), end="", flush=True)
        print("\rbatch:[%8d] %d sec elapsed" % (cnt, int(time.time() - start)))
        with open(self.name_pred + ".meta", "w") as fm:
            fm.write("%d,%d,%d,%d\n" % ((nrow,) + dshape))
        print("finished prediction, write into %s" % self.name_pred)

    def task_train(self) -> None:
        start = time.time()
        elapsed = 0
        err = sys.stderr
        if self.continue_training == 0 and self.name_model_in == "NULL":
            self.save_model()
        else:
            if not self.silent:
                print("continuing from round %d" % (self.start_counter - 1), end="", flush=True)
            for it, nm in zip(self.itr_evals, self.eval_names):
                res = self.net_trainer.evaluate(it, nm)
                if self.rank == 0:
                    err.write(res)
            if self.rank == 0:
                err.write("\n")
                err.flush()
        if self.itr_train is None:
            return
        if self.test_io != 0:
            print("start I/O test")
        cc = self.max_round
        while self.start_counter <= self.num_round and cc > 0:
            cc -= 1
            if not self.silent and self.rank == 0:
                print("update round %d" % (self.start_counter - 1), end="", flush=True)
            sample_counter = 0
            round_t0, n_inst = time.time(), 0
            self.net_trainer.start_round(self.start_counter)
            self.itr_train.before_first()
            while self.itr_train.next():
                if self.test_io == 0:
                    self.net_trainer.update(self.itr_train.value())
                else:
                    n_inst += self.itr_train.value().batch_size
                sample_counter += 1
                if sample_counter % self.print_step == 0:
                    elapsed = int(time.time() - start)
                    if not self.silent and self.rank == 0:
                        print("\r                                                               \r", end="")
                        print("round %8d:[%8d] %d sec elapsed" % (self.start_counter - 1, sample_counter, elapsed),
                              end="", flush=True)
            if self.test_io == 0:
                line = "[%d]" % self.start_counter
                if not self.itr_evals:
                    line += self.net_trainer.evaluate(None, "train")
                for it, nm in zip(self.itr_evals, self.eval_names):
                    line += self.net_trainer.evaluate(it, nm)
                if self.rank == 0:
                    err.write(line + "\n")
                    err.flush()
            if self.test_io != 0 and self.rank == 0:
                # the reference only runs the pipeline; the rate is what one wants to know from such a run
                dt = max(time.time() - round_t0, 1e-9)
                print("\nI/O test round %d: %d batches, %d instances in %.2f sec: %.0f instances/sec"
                      % (self.start_counter - 1, sample_counter, n_inst, dt, n_inst / dt))
            elapsed = int(time.time() - start)
            report = self.net_trainer.timing_report()
            if report and self.rank == 0:
                print("\n" + report)
            self.save_model()
        if not self.silent and self.rank == 0:
            print("\nupdating end, %d sec in all" % elapsed)


def main(argv: Optional[List[str]] = None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    return LearnTask().run(argv)


if __name__ == "__main__":
    sys.exit(main())
