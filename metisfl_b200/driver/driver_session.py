"""Driver: builds a federation from a YAML environment, launches it, monitors it, tears it down.

Same four-call session API as the reference (``metisfl/driver/driver_session.py:491-585``):
``DriverSession(fed_env, model, train/validation/test_dataset_recipe_fn)`` then
``initialize_federation()``, ``monitor_federation()``, ``shutdown_federation()``,
``get_federation_statistics()`` (the same four statistics sections, :395-418).

Launching differs by design.  The reference always goes through SSH (fabric) - even to localhost,
relying on an sshd with root/root in its Docker image (:506-582).  On a single 8xB200 box every
process is local, so the default launcher starts the controller and the learners as child
processes with ``CUDA_VISIBLE_DEVICES`` set per learner (``Launcher: local``); ``Launcher:
inprocess`` keeps everything inside the driver process (tests, CPU plumbing), and ``Launcher: ssh``
preserves the reference behaviour for off-box learners.
"""
from __future__ import annotations

import inspect
import os
import shlex
import shutil
import subprocess
import sys
import tarfile
import threading
import time

import cloudpickle
import torch
from google.protobuf.json_format import MessageToDict

from metisfl_b200.models.torch_model_ops import TorchModelOps, save_torch_model
from metisfl_b200.proto import metis_pb2
from metisfl_b200.utils import fedenv_parser, proto_messages_factory
from metisfl_b200.utils.grpc_controller_client import GRPCControllerClient
from metisfl_b200.utils.grpc_learner_client import GRPCLearnerClient
from metisfl_b200.utils.init_services_factory import MetisInitServicesCmdFactory
from metisfl_b200.utils.metis_logger import MetisASCIIArt, MetisLogger
from metisfl_b200.utils.ssl_configurator import SSLConfigurator

_P = proto_messages_factory.MetisProtoMessages
_M = proto_messages_factory.ModelProtoMessages


class DriverSessionBase(object):

    def __init__(self, fed_env, model, train_dataset_recipe_fn, validation_dataset_recipe_fn=None,
                 test_dataset_recipe_fn=None, working_dir="/tmp/metis/", launcher=None, model_dtype=None,
                 banner=True, neural_engine="pytorch", echo_sleep_secs=3.0):
        # neural_engine="echo": learners that sleep ``echo_sleep_secs`` and echo the model back (controller load test)
        self.neural_engine, self.echo_sleep_secs = neural_engine, echo_sleep_secs
        if banner:
            MetisASCIIArt.print()
        self.federation_environment = fed_env if isinstance(fed_env, fedenv_parser.FederationEnvironment) \
            else fedenv_parser.FederationEnvironment(fed_env)
        self.num_participating_learners = len(self.federation_environment.learners.learners)
        self.launcher = (launcher or self.federation_environment.launcher or "local").lower()
        self.working_dir = working_dir
        if os.path.exists(working_dir):
            shutil.rmtree(working_dir)
        os.makedirs(working_dir)

        if not isinstance(model, torch.nn.Module):
            raise RuntimeError("Not a supported model type (a torch.nn.Module is required).")
        self._save_model_dir_name = "model_definition"
        self._save_model_dir = os.path.join(working_dir, self._save_model_dir_name)
        try:  # ship user-defined model classes by value so learners need not import the user's module
            mod = inspect.getmodule(model)
            if mod is not None and not mod.__name__.startswith(("torch", "metisfl_b200")):
                cloudpickle.register_pickle_by_value(mod)
        except Exception:
            pass
        save_torch_model(model, self._save_model_dir)
        # names / trainable flags / values of the FULL state in the dtype the learners will use
        self._model_dtype = model_dtype
        self._model_weights_descriptor = self._describe_model(model)
        self.model_definition_tar_fp = self._make_tarfile(self._save_model_dir_name, self._save_model_dir)

        def _dump(fn, name):
            if not fn:
                return None
            path = os.path.join(working_dir, name)
            with open(path, "wb+") as f:
                cloudpickle.dump(obj=fn, file=f)
            return path

        if not train_dataset_recipe_fn:
            raise RuntimeError("Train dataset recipe cannot be empty.")
        self.train_dataset_recipe_fp = _dump(train_dataset_recipe_fn, "model_train_dataset_ops.pkl")
        self.validation_dataset_recipe_fp = _dump(validation_dataset_recipe_fn, "model_validation_dataset_ops.pkl")
        self.test_dataset_recipe_fp = _dump(test_dataset_recipe_fn, "model_test_dataset_ops.pkl")

        self.enable_ssl = self.federation_environment.communication_protocol.enable_ssl
        self._children = []          # (name, Popen)
        self._inprocess = []         # in-process service objects
        self._driver_controller_grpc_client = self._create_driver_controller_grpc_client()
        self._driver_learner_grpc_clients = self._create_driver_learner_grpc_clients()
        self._federation_statistics = dict()
        self._init_homomorphic_encryption()
        self._fabric_specs = self._plan_device_fabric()

    # -- device fabric ------------------------------------------------------------------------------
    def _plan_device_fabric(self):
        """Decide whether the learners form an on-box NVLink fabric (models stay in HBM, messages carry device-slot
        references) and give every learner its rank.  Returns ``{learner index: "rank=..,world=..,..."}``."""
        env = self.federation_environment
        want = env.device_fabric
        want = str(want).lower() if not isinstance(want, bool) else want
        if want in (False, "false", "off", "no", "0") or self.neural_engine == "echo" or self.launcher == "ssh":
            return {}
        rule = env.global_model_config.aggregation_rule.aggregation_rule_name
        members = [i for i, l in enumerate(env.learners) if getattr(l, "in_fabric", True)]
        n = len(members)
        forced = want in (True, "true", "on", "yes", "1")
        ok = n > 0 and rule in ("FedAvg", "FedStride", "FedRec", "FedAsync") and self._he_scheme is None
        # encrypted FedAvg on the fabric (ciphertexts in peer-visible HBM, K8 reduces them over NVLink): on request
        # only (`DeviceFabric: true`), and only when every learner is a member - a ciphertext that arrives as bytes
        # cannot be added to one that lives on a GPU
        if rule == "PWA" and self._he_scheme is not None and forced and n == len(env.learners) and n > 0:
            ok = True
        # the asynchronous rules keep their rolling state on the executor GPU: a learner outside the fabric would be
        # folded on the host into a second, diverging community model - such a federation stays on the host path
        ok = ok and not (rule in ("FedRec", "FedAsync") and n < len(env.learners))
        ok = ok and (self.launcher == "local" or n == 1)   # one process is one rank: in-process only with one member
        if not forced:
            ok = ok and torch.cuda.is_available() and n <= torch.cuda.device_count() and n <= 16
        if not ok:
            if forced:
                raise RuntimeError("DeviceFabric was requested but rule {} / launcher {} / {} learners cannot run on "
                                   "one fabric".format(rule, self.launcher, n))
            return {}
        import socket
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
        s.close()
        dtype = env.model_dtype or (self._model_dtype and {torch.bfloat16: "bf16", torch.float32: "fp32",
                                                           torch.float16: "fp16"}.get(self._model_dtype))
        specs = {}
        for rank, i in enumerate(members):   # learners outside the fabric (`DeviceFabric: false`) keep the host path
            t = "rank={},world={},init=tcp://127.0.0.1:{},id=fabric0,rule={},staging=2,executor=0".format(rank, n, port, rule)
            specs[i] = t + (",dtype={}".format(dtype) if dtype else "")
        MetisLogger.info("Device fabric: %d learner(s), rule %s, models stay on the GPUs", n, rule)
        return specs

    # -- helpers ----------------------------------------------------------------------------------
    def _describe_model(self, model):
        import copy
        ops = TorchModelOps(model=copy.deepcopy(model), device="cpu",
                            model_dtype=self._model_dtype or torch.float32, use_cuda_graph=False)
        return ops.get_model_weights()

    def _make_tarfile(self, output_filename, source_dir):
        output_dir = os.path.abspath(os.path.join(source_dir, os.pardir))
        output_filepath = os.path.join(output_dir, "{}.tar.gz".format(output_filename))
        with tarfile.open(output_filepath, "w:gz") as tar:
            tar.add(source_dir, arcname=os.path.basename(source_dir))
        return output_filepath

    def _init_homomorphic_encryption(self):
        env = self.federation_environment
        self._he_scheme = None
        self._crypto_params_dir = os.path.join(self.working_dir, "cryptoparams")
        os.makedirs(self._crypto_params_dir, exist_ok=True)
        if env.homomorphic_encryption and (env.homomorphic_encryption.scheme or "").upper() == "CKKS":
            from metisfl_b200.encryption import fhe
            he = env.homomorphic_encryption
            self._he_scheme = fhe.CKKS(he.batch_size, he.scaling_factor_bits)
            self._he_scheme.gen_crypto_context_and_keys(self._crypto_params_dir)
            files = self._crypto_params_files = self._he_scheme.get_crypto_params_files()
            ckks_pb = _P.construct_ckks_scheme_config_pb(batch_size=he.batch_size,
                                                         scaling_factor_bits=he.scaling_factor_bits)
            # controller: crypto context only; learners: context + both keys
            self._controller_he_scheme_config_pb = _P.construct_he_scheme_config_pb(
                enabled=True, crypto_context_file=files["crypto_context_file"], ckks_scheme_config_pb=ckks_pb)
            self._learners_he_scheme_config_pb = _P.construct_he_scheme_config_pb(
                enabled=True, crypto_context_file=files["crypto_context_file"],
                public_key_file=files["public_key_file"], private_key_file=files["private_key_file"],
                ckks_scheme_config_pb=ckks_pb)
        else:
            empty = _P.construct_empty_scheme_config_pb()
            self._controller_he_scheme_config_pb = _P.construct_he_scheme_config_pb(
                enabled=False, empty_scheme_config_pb=empty)
            self._learners_he_scheme_config_pb = _P.construct_he_scheme_config_pb(
                enabled=False, empty_scheme_config_pb=empty)

    def _ssl_config_pb(self, remote):
        if not self.enable_ssl:
            return _P.construct_ssl_config_pb(enable_ssl=False)
        cert = key = None
        if remote.ssl_configs is not None:
            cert, key = remote.ssl_configs.public_certificate_file, remote.ssl_configs.private_key_file
        if not (cert and key):
            cert, key = SSLConfigurator.gen_default_certificates(as_stream=False)
        return _P.construct_ssl_config_pb(enable_ssl=True, config_pb=_P.construct_ssl_config_files_pb(cert, key))

    def _server_entity_pb(self, remote):
        return _P.construct_server_entity_pb(hostname=remote.grpc_servicer.hostname, port=remote.grpc_servicer.port,
                                             ssl_config_pb=self._ssl_config_pb(remote))

    def _create_driver_controller_grpc_client(self):
        return GRPCControllerClient(self._server_entity_pb(self.federation_environment.controller), max_workers=1)

    def _create_driver_learner_grpc_clients(self):
        return {l.learner_id or "{}:{}".format(l.grpc_servicer.hostname, l.grpc_servicer.port):
                GRPCLearnerClient(self._server_entity_pb(l), max_workers=1)
                for l in self.federation_environment.learners}

    # -- protobuf configuration -------------------------------------------------------------------
    def _controller_params_pb(self):
        env = self.federation_environment
        rule = env.global_model_config.aggregation_rule
        aggregation_rule_pb = _P.construct_aggregation_rule_pb(
            rule_name=rule.aggregation_rule_name, scaling_factor=rule.aggregation_rule_scaling_factor,
            stride_length=rule.aggregation_rule_stride_length,
            he_scheme_config_pb=self._controller_he_scheme_config_pb, fed_async_kwargs=rule.fed_async_kwargs)
        global_model_specs_pb = _P.construct_global_model_specs(
            aggregation_rule_pb=aggregation_rule_pb,
            learners_participation_ratio=env.global_model_config.participation_ratio)
        proto = env.communication_protocol
        communication_specs_pb = _P.construct_communication_specs_pb(
            protocol=proto.name, semi_sync_lambda=proto.semi_synchronous_lambda,
            semi_sync_recompute_num_updates=proto.semi_sync_recompute_num_updates,
            task_deadline_secs=proto.task_deadline_secs)
        optimizer_pb = _M.construct_optimizer_config_pb_from_kwargs(env.local_model_config.optimizer_config.optimizer_pb_kwargs)
        hyper_pb = _P.construct_controller_modelhyperparams_pb(
            batch_size=env.local_model_config.batch_size, epochs=env.local_model_config.local_epochs,
            optimizer_pb=optimizer_pb, percent_validation=env.local_model_config.validation_percentage)
        store = env.model_store_config
        store_pb = _P.construct_model_store_config_pb(
            name=store.name, eviction_policy=store.eviction_policy, lineage_length=store.eviction_lineage_length,
            store_hostname=store.connection_configs.hostname, store_port=store.connection_configs.port)
        return _P.construct_controller_params_pb(self._server_entity_pb(env.controller), global_model_specs_pb,
                                                 communication_specs_pb, store_pb, hyper_pb)

    def _init_controller_cmd(self):
        p = self._controller_params_pb()
        return MetisInitServicesCmdFactory.init_controller_target(
            controller_server_entity_pb_ser=p.server_entity.SerializeToString(),
            global_model_specs_pb_ser=p.global_model_specs.SerializeToString(),
            communication_specs_pb_ser=p.communication_specs.SerializeToString(),
            model_hyperparameters_pb_ser=p.model_hyperparams.SerializeToString(),
            model_store_config_pb_ser=p.model_store_config.SerializeToString())

    def _init_learner_cmd(self, learner, model_dir, recipes, device, fabric=None):
        return MetisInitServicesCmdFactory.init_learner_target(
            learner_server_entity_pb_ser=self._server_entity_pb(learner).SerializeToString(),
            controller_server_entity_pb_ser=self._server_entity_pb(self.federation_environment.controller).SerializeToString(),
            he_scheme_pb_ser=self._learners_he_scheme_config_pb.SerializeToString(),
            model_dir=model_dir, train_dataset=learner.dataset_configs.train_dataset_path or "",
            validation_dataset=learner.dataset_configs.validation_dataset_path,
            test_dataset=learner.dataset_configs.test_dataset_path,
            train_dataset_recipe=recipes[0], validation_dataset_recipe=recipes[1] or "",
            test_dataset_recipe=recipes[2] or "", neural_engine=self.neural_engine, device=device,
            fabric=fabric) + (
                " --echo_sleep_secs={}".format(self.echo_sleep_secs) if self.neural_engine == "echo" else "")

    # -- launchers --------------------------------------------------------------------------------
    def _popen(self, name, cmd, extra_env=None):
        env = dict(os.environ)
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
        env.update(extra_env or {})
        log = open(os.path.join(self.working_dir, f"{name}.log"), "wb")
        proc = subprocess.Popen(shlex.split(cmd), env=env, stdout=log, stderr=subprocess.STDOUT,
                                start_new_session=True)
        self._children.append((name, proc))
        return proc

    def _init_controller(self):
        if self.launcher == "inprocess":
            from metisfl_b200.controller.controller_instance import ControllerInstance
            inst = ControllerInstance()
            inst.start(self._controller_params_pb())
            self._inprocess.append(("controller", inst))
        elif self.launcher == "ssh":
            self._ssh_run(self.federation_environment.controller, self._init_controller_cmd(), files=[])
        else:
            self._popen("controller", self._init_controller_cmd())

    def _init_learner(self, index, learner):
        recipes = (self.train_dataset_recipe_fp, self.validation_dataset_recipe_fp, self.test_dataset_recipe_fp)
        cuda = ",".join(str(d) for d in (learner.cuda_devices or []))
        device = None
        if not torch.cuda.is_available() or not learner.cuda_devices:
            device = "cpu" if not torch.cuda.is_available() else None
        if self.launcher == "inprocess":
            from metisfl_b200.learner.learner import Learner
            from metisfl_b200.learner.learner_servicer import LearnerServicer
            dev = device or ("cuda:{}".format(learner.cuda_devices[0]) if learner.cuda_devices else None)
            if self.neural_engine == "echo":
                from metisfl_b200.learner.echo_learner import EchoLearner
                lrn = EchoLearner(self._server_entity_pb(learner),
                                  self._server_entity_pb(self.federation_environment.controller),
                                  metis_pb2.HESchemeConfig(), sleep_secs=self.echo_sleep_secs,
                                  learner_credentials_fp=os.path.join(self.working_dir, "learner_credentials"))
            else:
                lrn = Learner(self._server_entity_pb(learner), self._server_entity_pb(self.federation_environment.controller),
                              self._learners_he_scheme_config_pb, "pytorch", self._save_model_dir,
                              learner.dataset_configs.train_dataset_path, recipes[0],
                              learner.dataset_configs.validation_dataset_path, recipes[1] or "",
                              learner.dataset_configs.test_dataset_path, recipes[2] or "", device=dev,
                              learner_credentials_fp=os.path.join(self.working_dir, "learner_credentials"),
                              fabric=self._fabric_specs.get(index))
            svc = LearnerServicer(lrn, servicer_workers=5)
            svc.init_servicer()
            t = threading.Thread(target=svc.wait_servicer, daemon=True)
            t.start()
            self._inprocess.append((f"learner{index}", svc))
        elif self.launcher == "ssh":
            remote_dir = "/tmp/metis/workdir_learner_{}".format(learner.grpc_servicer.port)
            files = [self.model_definition_tar_fp] + [r for r in recipes if r]
            rr = [os.path.join(remote_dir, os.path.basename(r)) if r else None for r in recipes]
            cmd = "mkdir -p {d} && cd {d} && tar -xzf {t} && ".format(
                d=remote_dir, t=os.path.basename(self.model_definition_tar_fp))
            if cuda:
                cmd += "export CUDA_VISIBLE_DEVICES={} && ".format(cuda)
            cmd += "cd {} && ".format(learner.project_home) if learner.project_home else ""
            cmd += self._init_learner_cmd(learner, os.path.join(remote_dir, self._save_model_dir_name), rr, device)
            self._ssh_run(learner, cmd, files=files, remote_dir=remote_dir)
        else:
            fabric = self._fabric_specs.get(index)
            if fabric and torch.cuda.is_available():
                # peers map each other's memory: every GPU stays visible and the learner is told which one is its own
                env = {}
                device = "cuda:{}".format(learner.cuda_devices[0] if learner.cuda_devices else index)
            else:
                env = {"CUDA_VISIBLE_DEVICES": cuda} if cuda else {}
            self._popen(f"learner{index}", self._init_learner_cmd(learner, self._save_model_dir, recipes, device,
                                                                 fabric=fabric), env)

    def _ssh_run(self, remote, cmd, files, remote_dir=None):
        """Reference-style remote launch over SSH/SFTP (fabric); runs in a background thread."""
        import fabric

        def _go():
            conn = fabric.Connection(**remote.connection_configs.get_fabric_connection_config())
            if remote_dir:
                conn.run("mkdir -p {}".format(remote_dir))
                for f in files:
                    conn.put(f, remote_dir)
            prefix = remote.connection_configs.on_login
            full = "{} {}".format(prefix, cmd) if prefix and prefix != "clear" else cmd
            if remote.project_home and not remote_dir:
                full = "cd {} && {}".format(remote.project_home, full)
            conn.run(full)
            conn.close()
        t = threading.Thread(target=_go, daemon=True)
        t.start()
        self._children.append((f"ssh:{remote.grpc_servicer.port}", t))

    # -- lifecycle --------------------------------------------------------------------------------
    def _ship_model_to_controller(self):
        names, trainables, values = self._model_weights_descriptor
        model_pb = _M.construct_model_pb_from_np(values, names, trainables, he_scheme=self._he_scheme)
        self._driver_controller_grpc_client.replace_community_model(num_contributors=self.num_participating_learners,
                                                                    model_pb=model_pb)

    def initialize_federation(self, model_weights=None):
        if model_weights is not None:
            self._model_weights_descriptor = model_weights
        self._init_controller()
        self._driver_controller_grpc_client.check_health_status(request_retries=30, request_timeout=30, block=True)
        self._ship_model_to_controller()
        for index, learner in enumerate(self.federation_environment.learners):
            self._init_learner(index, learner)

    def _collect_local_statistics(self):
        learners_pb = self._driver_controller_grpc_client.get_participating_learners()
        learners_id = [l.id for l in learners_pb.learner]
        self._federation_statistics["learners_descriptor"] = MessageToDict(learners_pb, preserving_proto_field_name=True)
        results = self._driver_controller_grpc_client.get_local_task_lineage(-1, learners_id)
        self._federation_statistics["learners_models_results"] = MessageToDict(results, preserving_proto_field_name=True)

    def _collect_global_statistics(self):
        md = self._driver_controller_grpc_client.get_runtime_metadata(num_backtracks=0)
        self._federation_statistics["federation_runtime_metadata"] = MessageToDict(md, preserving_proto_field_name=True)
        ev = self._driver_controller_grpc_client.get_community_model_evaluation_lineage(-1)
        self._federation_statistics["community_model_results"] = MessageToDict(ev, preserving_proto_field_name=True)

    def get_federation_statistics(self):
        return self._federation_statistics

    def _children_alive(self):
        for name, proc in self._children:
            if isinstance(proc, subprocess.Popen) and proc.poll() is not None and proc.returncode != 0:
                raise RuntimeError("{} exited with code {} (see {}/{}.log)".format(
                    name, proc.returncode, self.working_dir, name))

    def monitor_federation(self, request_every_secs=10):
        sig = self.federation_environment.termination_signals
        proto = self.federation_environment.communication_protocol
        metric = self.federation_environment.evaluation_metric
        start = time.time()
        evals_checked_up_to = 0
        while True:
            time.sleep(request_every_secs)
            self._children_alive()
            # the newest entries only: the reference asks for the whole lineage on every poll (driver_session.py:
            # 423-480), which grows with the number of rounds and is served under the controller's lock
            metadata = self._driver_controller_grpc_client.get_runtime_metadata(num_backtracks=1).metadata
            if (proto.is_synchronous or proto.is_semi_synchronous) and sig.federation_rounds and len(metadata) > 0:
                if max(m.global_iteration for m in metadata) > sig.federation_rounds:
                    MetisLogger.info("Exceeded federation rounds cutoff point. Exiting ...")
                    return
            if proto.is_asynchronous and sig.federation_rounds and len(metadata) > 0:
                # asynchronous: count community updates, one per learner contribution
                if max(m.global_iteration for m in metadata) > sig.federation_rounds * max(1, self.num_participating_learners):
                    MetisLogger.info("Exceeded community update cutoff point. Exiting ...")
                    return
            current = max((m.global_iteration for m in metadata), default=0)
            # evaluations filed since the last poll, plus the two newest ones again (their answers may still be arriving)
            window = max(4, current - evals_checked_up_to + 2)
            results = self._driver_controller_grpc_client.get_community_model_evaluation_lineage(window)
            evals_checked_up_to = max(0, current - 2)
            for res in results.community_evaluation:
                scores = [float(ev.test_evaluation.metric_values[metric]) for ev in res.evaluations.values()
                          if metric in ev.test_evaluation.metric_values]
                if scores and sum(scores) / len(scores) >= sig.metric_cutoff_score:
                    MetisLogger.info("Exceeded evaluation metric cutoff score. Exiting ...")
                    return
            if (time.time() - start) / 60.0 > sig.execution_time_cutoff_mins:
                MetisLogger.info("Exceeded execution time cutoff minutes. Exiting ...")
                return

    def _shutdown(self):
        try:
            self._collect_local_statistics()
        except Exception as e:
            MetisLogger.warning("could not collect local statistics: %r", e)
        for lid, client in self._driver_learner_grpc_clients.items():
            try:
                client.shutdown_learner(request_retries=1, request_timeout=30, block=True)
            except Exception as e:
                MetisLogger.warning("learner %s shutdown failed: %r", lid, e)
        time.sleep(0.5)  # learners leave the federation before the controller stops
        try:
            self._collect_global_statistics()
        except Exception as e:
            MetisLogger.warning("could not collect global statistics: %r", e)
        try:
            self._driver_controller_grpc_client.shutdown_controller(request_retries=2, request_timeout=30, block=True)
        except Exception as e:
            MetisLogger.warning("controller shutdown failed: %r", e)
        for client in self._driver_learner_grpc_clients.values():
            client.shutdown()
        self._driver_controller_grpc_client.shutdown()
        for name, inst in self._inprocess:
            if name == "controller":
                inst.shutdown(instantly=True)
        deadline = time.time() + 20
        for name, proc in self._children:
            if isinstance(proc, subprocess.Popen):
                try:
                    proc.wait(timeout=max(0.1, deadline - time.time()))
                except subprocess.TimeoutExpired:
                    proc.terminate()  # exactly the process we started
                    try:
                        proc.wait(timeout=5)
                    except subprocess.TimeoutExpired:
                        proc.kill()


class DriverSession(DriverSessionBase):

    def shutdown_federation(self):
        self._shutdown()

    def run(self, request_every_secs=2):
        """Convenience: initialize, monitor until a termination signal fires, shut down, return statistics."""
        self.initialize_federation()
        try:
            self.monitor_federation(request_every_secs=request_every_secs)
        finally:
            self.shutdown_federation()
        return self.get_federation_statistics()
