"""YAML federation environment -> Python objects.

Same schema and attribute names as the reference parser (``metisfl/utils/fedenv_parser.py:10-310``;
schema summarised in SURVEY §5.6) so existing environment files load unchanged.  Implemented as a
small declarative mapping (YAML key -> attribute, default) instead of one hand-written class body per
section.  Extensions: ``FedAsync`` rule parameters, ``TaskDeadlineSecs``, a ``Launcher`` switch
(``local`` subprocesses by default; ``ssh`` for the reference's fabric behaviour) and per-learner
``LogicalLearners`` for multiplexing.
"""
from __future__ import annotations

import yaml

_REQUIRED = object()


class _Section(object):
    """``FIELDS`` = (attribute, yaml key, default | _REQUIRED | callable(section_map) )."""
    FIELDS = ()

    def __init__(self, section_map):
        section_map = section_map or {}
        self._map = section_map
        for attr, key, default in self.FIELDS:
            if key in section_map and section_map[key] is not None:
                value = section_map[key]
            elif default is _REQUIRED:
                raise AssertionError(f"{type(self).__name__}: missing required key '{key}'")
            else:
                value = default
            setattr(self, attr, value)

    def __repr__(self):
        body = ", ".join(f"{a}={getattr(self, a)!r}" for a, _, _ in self.FIELDS)
        return f"{type(self).__name__}({body})"


class DockerImage(object):
    def __init__(self, docker_image):
        self.docker_image = docker_image


class TerminationSignals(_Section):
    FIELDS = (("federation_rounds", "FederationRounds", 100),
              ("execution_time_cutoff_mins", "ExecutionCutoffTimeMins", 1e6),
              ("metric_cutoff_score", "MetricCutoffScore", 1))


class HomomorphicEncryption(_Section):
    FIELDS = (("scheme", "Scheme", None), ("batch_size", "BatchSize", 4096),
              ("scaling_factor_bits", "ScalingFactorBits", 52))

    def __init__(self, m):
        super().__init__(m)
        if self.scheme is not None and self.scheme.upper() != "CKKS":
            raise AssertionError("Only the CKKS homomorphic encryption scheme is supported.")


class CommunicationProtocol(_Section):
    FIELDS = (("enable_ssl", "EnableSSL", False), ("name", "Name", _REQUIRED), ("specifications", "Specifications", None))

    def __init__(self, m):
        super().__init__(m)
        self.enable_ssl = bool(self.enable_ssl)
        key = self.name.upper().replace("_", "")
        self.is_asynchronous = key == "ASYNCHRONOUS"
        self.is_synchronous = key == "SYNCHRONOUS"
        self.is_semi_synchronous = key == "SEMISYNCHRONOUS"
        spec = self.specifications or {}
        self.semi_synchronous_lambda = spec.get("SemiSynchronousLambda", None)
        self.semi_sync_recompute_num_updates = spec.get("SemiSynchronousRecomputeSteps", None)
        self.task_deadline_secs = spec.get("TaskDeadlineSecs", None)


class AggregationRule(_Section):
    FIELDS = (("aggregation_rule_name", "Name", None), ("aggregation_rule_specifications", "RuleSpecifications", {}))

    def __init__(self, m):
        super().__init__(m)
        spec = self.aggregation_rule_specifications or {}
        self.aggregation_rule_scaling_factor = spec.get("ScalingFactor", None)
        self.aggregation_rule_stride_length = spec.get("StrideLength", None)
        self.fed_async_kwargs = {k: spec[y] for k, y in (("alpha", "Alpha"), ("staleness_fn", "StalenessFunction"),
                                                         ("a", "StalenessA"), ("b", "StalenessB")) if y in spec}

    def __str__(self):
        return "AggregationRule(name={}, scaling_factor={}, stride_length={})".format(
            self.aggregation_rule_name, self.aggregation_rule_scaling_factor, self.aggregation_rule_stride_length)


class GlobalModelConfig(object):
    def __init__(self, global_model_map):
        global_model_map = global_model_map or {}
        self.aggregation_rule = AggregationRule(global_model_map.get("AggregationRule", None))
        self.participation_ratio = global_model_map.get("ParticipationRatio", 1)


class OptimizerConfig(object):
    _KW = {
        "VANILLASGD": ("VanillaSGD", {"L1Reg": "l1_reg", "L2Reg": "l2_reg"}, {}),
        "MOMENTUMSGD": ("MomentumSGD", {"MomentumFactor": "momentum_factor"}, {}),
        "FEDPROX": ("FedProx", {"ProximalTerm": "proximal_term"}, {}),
        "ADAM": ("Adam", {"Beta1": "beta_1", "Beta2": "beta_2", "Epsilon": "epsilon"}, {}),
        "ADAMW": ("AdamWeightDecay", {"WeightDecay": "weight_decay"}, {"weight_decay": 1e-4}),
    }

    def __init__(self, optimizer_map):
        optimizer_map = optimizer_map or {}
        self.optimizer_name = optimizer_map.get("OptimizerName")
        self.learning_rate = optimizer_map.get("LearningRate")
        self.optimizer_pb_kwargs = self.create_optimizer_pb_kwargs(optimizer_map)

    def create_optimizer_pb_kwargs(self, optimizer_map):
        try:
            pb_name, keys, defaults = self._KW[str(self.optimizer_name).upper()]
        except KeyError:
            raise RuntimeError("Not a supported optimizer.") from None
        kwargs = {"learning_rate": self.learning_rate, "name": pb_name, **defaults}
        kwargs.update({kw: optimizer_map[y] for y, kw in keys.items() if y in optimizer_map})
        return kwargs


class LocalModelConfig(object):
    def __init__(self, local_model_map):
        local_model_map = local_model_map or {}
        self.batch_size = local_model_map.get("BatchSize", 100)
        self.local_epochs = local_model_map.get("LocalEpochs", 5)
        self.validation_percentage = local_model_map.get("ValidationPercentage", 0)
        self.optimizer_config = OptimizerConfig(local_model_map.get("OptimizerConfig"))


class ConnectionConfigsBase(_Section):
    FIELDS = (("hostname", "Hostname", None), ("port", "Port", None))


class ConnectionConfigs(_Section):
    FIELDS = (("hostname", "Hostname", None), ("port", "Port", None), ("username", "Username", ""),
              ("password", "Password", ""), ("key_filename", "KeyFilename", ""), ("passphrase", "Passphrase", ""),
              ("on_login", "OnLogin", "clear"))

    def get_fabric_connection_config(self):
        """kwargs for ``fabric.Connection`` (password or key-file authentication)."""
        connect_kwargs = {"password": self.password, "allow_agent": not self.password,
                          "look_for_keys": bool(self.key_filename)}
        if self.key_filename:
            connect_kwargs.update(key_filename=self.key_filename, passphrase=self.passphrase)
        return {"host": self.hostname, "port": self.port, "user": self.username, "connect_kwargs": connect_kwargs}


class ModelStoreConfig(object):
    def __init__(self, model_store_map):
        if model_store_map is None:  # default when the YAML omits the section
            model_store_map = {"Name": "InMemory", "EvictionPolicy": "LineageLengthEviction", "LineageLength": 1}
        self.name = model_store_map.get("Name", None)
        self.eviction_policy = model_store_map.get("EvictionPolicy")
        self.eviction_lineage_length = model_store_map.get("LineageLength", 1)
        self.connection_configs = ConnectionConfigsBase(model_store_map.get("ConnectionConfigs", {}))


class SSLConfigs(_Section):
    FIELDS = (("public_certificate_file", "PublicCertificate", None), ("private_key_file", "PrivateKey", None))


class GRPCServicer(_Section):
    FIELDS = (("hostname", "Hostname", None), ("port", "Port", None),
              ("public_certificate_path", "PublicCertificatePath", None),
              ("private_key_path", "PrivateKeyPath", None))


class DatasetConfigs(_Section):
    FIELDS = (("train_dataset_path", "TrainDatasetPath", None), ("validation_dataset_path", "ValidationDatasetPath", ""),
              ("test_dataset_path", "TestDatasetPath", ""))


class RemoteHost(object):
    def __init__(self, config_map):
        config_map = config_map or {}
        self.connection_configs = ConnectionConfigs(config_map.get("ConnectionConfigs"))
        self.grpc_servicer = GRPCServicer(config_map.get("GRPCServicer"))
        self.ssl_configs = SSLConfigs(config_map.get("SSLConfigs")) if "SSLConfigs" in config_map else None
        self.project_home = config_map.get("ProjectHome", "")


class Controller(RemoteHost):
    pass


class Learner(RemoteHost):
    def __init__(self, learner_def_map):
        super().__init__(learner_def_map)
        self.learner_id = learner_def_map.get("LearnerID")
        self.cuda_devices = learner_def_map.get("CudaDevices", [])
        self.logical_learners = learner_def_map.get("LogicalLearners", 1)
        # extension: `DeviceFabric: false` keeps THIS learner on the host path (model bytes through the controller) even
        # when the rest of the federation forms an on-box fabric - e.g. a learner on another machine
        self.in_fabric = str(learner_def_map.get("DeviceFabric", True)).lower() not in ("false", "off", "no", "0")
        self.dataset_configs = DatasetConfigs(learner_def_map.get("DatasetConfigs"))

    def __str__(self):
        return "Learner(id={}, grpc={}:{}, cuda={})".format(
            self.learner_id, self.grpc_servicer.hostname, self.grpc_servicer.port, self.cuda_devices)


class Learners(object):
    def __init__(self, learners_map):
        self.learners = [Learner(d) for d in (learners_map or [])]

    def __iter__(self):
        return iter(self.learners)

    def __len__(self):
        return len(self.learners)

    def __getitem__(self, i):
        return self.learners[i]


class FederationEnvironment(object):

    def __init__(self, federation_environment_config_fp):
        if isinstance(federation_environment_config_fp, dict):
            self.loaded_stream = federation_environment_config_fp
        else:
            with open(federation_environment_config_fp) as fstream:
                self.loaded_stream = yaml.load(fstream, Loader=yaml.SafeLoader)
        env = self.loaded_stream.get("FederationEnvironment")
        self.docker = DockerImage(env.get("DockerImage"))
        self.termination_signals = TerminationSignals(env.get("TerminationSignals"))
        self.evaluation_metric = env.get("EvaluationMetric")
        self.communication_protocol = CommunicationProtocol(env.get("CommunicationProtocol"))
        self.global_model_config = GlobalModelConfig(env.get("GlobalModelConfig"))
        self.local_model_config = LocalModelConfig(env.get("LocalModelConfig"))
        self.model_store_config = ModelStoreConfig(env.get("ModelStoreConfig", None))
        self.controller = Controller(env.get("Controller"))
        self.learners = Learners(env.get("Learners"))
        self.launcher = str(env.get("Launcher", "local")).lower()
        # DeviceFabric (extension): learners of this box form one NVLink peer-memory fabric and exchange device-slot
        # references instead of model bytes.  "auto" = when CUDA is there, the rule has a device kernel and every
        # learner is local (the driver decides); true / false force it.
        self.device_fabric = env.get("DeviceFabric", "auto")
        self.model_dtype = env.get("ModelDType")  # bf16 | fp32 | fp16; default: bf16 on CUDA, fp32 on CPU
        self.homomorphic_encryption = None
        if "HomomorphicEncryption" in env and env.get("HomomorphicEncryption"):
            self.homomorphic_encryption = HomomorphicEncryption(env.get("HomomorphicEncryption"))
            # encrypted models can only be combined by the private weighted average
            assert self.global_model_config.aggregation_rule.aggregation_rule_name == "PWA", \
                "Since you have enabled Homomorphic Encryption, you need to use the PWA aggregation function."
