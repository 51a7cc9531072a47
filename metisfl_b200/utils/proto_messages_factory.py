"""Factories for the wire messages + the tensor <-> ``TensorSpec`` codec.

API parity with the reference's ``metisfl/utils/proto_messages_factory.py`` (class and method names
of ``ControllerServiceProtoMessages`` :9, ``LearnerServiceProtoMessages`` :49,
``MetisProtoMessages`` :89, ``ModelProtoMessages`` :397, ``ServiceCommonProtoMessages`` :636) so
code written against MetisFL keeps working.  The implementation is table-driven, adds torch
tensors, FLOAT16/BFLOAT16 and ``FedAsync``, and drops the reference's debug side effects
(``/tmp/test.npy`` dump and throw-away decrypt, :547-556).
"""
from __future__ import annotations

import sys

import numpy as np

from metisfl_b200.proto import (controller_pb2, learner_pb2, metis_pb2, model_pb2,
                                service_common_pb2)
from metisfl_b200.utils import dtypes

try:  # torch is optional for pure-protocol users
    import torch
except Exception:  # pragma: no cover
    torch = None


def _norm(name: str) -> str:
    return str(name).upper().replace("_", "").replace("-", "")


def _wrap_oneof(container_cls, message, **extra):
    """Put ``message`` into whichever oneof field of ``container_cls`` has its type."""
    for fd in container_cls.DESCRIPTOR.fields:
        if fd.message_type is not None and fd.message_type is message.DESCRIPTOR:
            return container_cls(**{fd.name: message}, **extra)
    raise RuntimeError(f"{type(message).__name__} is not a member of {container_cls.__name__}")


class ControllerServiceProtoMessages(object):

    @classmethod
    def construct_get_community_model_evaluation_lineage_request_pb(cls, num_backtracks):
        return controller_pb2.GetCommunityModelEvaluationLineageRequest(num_backtracks=num_backtracks)

    @classmethod
    def construct_get_community_model_lineage_request_pb(cls, num_backtracks):
        return controller_pb2.GetCommunityModelLineageRequest(num_backtracks=num_backtracks)

    @classmethod
    def construct_get_local_task_lineage_request_pb(cls, num_backtracks, learner_ids):
        return controller_pb2.GetLocalTaskLineageRequest(num_backtracks=num_backtracks, learner_ids=learner_ids)

    @classmethod
    def construct_get_runtime_metadata_lineage_request_pb(cls, num_backtracks):
        return controller_pb2.GetRuntimeMetadataLineageRequest(num_backtracks=num_backtracks)

    @classmethod
    def construct_get_participating_learners_request_pb(cls):
        return controller_pb2.GetParticipatingLearnersRequest()

    @classmethod
    def construct_join_federation_request_pb(cls, server_entity_pb, local_dataset_spec_pb):
        return controller_pb2.JoinFederationRequest(server_entity=server_entity_pb,
                                                    local_dataset_spec=local_dataset_spec_pb)

    @classmethod
    def construct_leave_federation_request_pb(cls, learner_id, auth_token):
        return controller_pb2.LeaveFederationRequest(learner_id=learner_id, auth_token=auth_token)

    @classmethod
    def construct_mark_task_completed_request_pb(cls, learner_id, auth_token, completed_learning_task_pb):
        return controller_pb2.MarkTaskCompletedRequest(learner_id=learner_id, auth_token=auth_token,
                                                       task=completed_learning_task_pb)

    @classmethod
    def construct_replace_community_model_request_pb(cls, federated_model_pb):
        return controller_pb2.ReplaceCommunityModelRequest(model=federated_model_pb)


class LearnerServiceProtoMessages(object):

    @classmethod
    def construct_evaluate_model_request_pb(cls, model=None, batch_size=None, eval_train=None,
                                            eval_test=None, eval_valid=None, metrics_pb=None):
        E = learner_pb2.EvaluateModelRequest
        which = [code for flag, code in ((eval_train, E.TRAINING), (eval_test, E.TEST), (eval_valid, E.VALIDATION))
                 if flag]
        return E(model=model, batch_size=batch_size, evaluation_dataset=which, metrics=metrics_pb)

    @classmethod
    def construct_evaluate_model_response_pb(cls, evaluation_pb=None):
        return learner_pb2.EvaluateModelResponse(evaluations=evaluation_pb)

    @classmethod
    def construct_run_task_request_pb(cls, federated_model_pb=None, learning_task_pb=None, hyperparameters_pb=None):
        return learner_pb2.RunTaskRequest(federated_model=federated_model_pb, task=learning_task_pb,
                                          hyperparameters=hyperparameters_pb)

    @classmethod
    def construct_run_task_response_pb(cls, ack_pb=None):
        return learner_pb2.RunTaskResponse(ack=ack_pb)


class MetisProtoMessages(object):
    _SCALING = {"NUMCOMPLETEDBATCHES": "NUM_COMPLETED_BATCHES", "NUMPARTICIPANTS": "NUM_PARTICIPANTS",
                "NUMTRAININGEXAMPLES": "NUM_TRAINING_EXAMPLES"}
    _PROTOCOL = {"SYNCHRONOUS": "SYNCHRONOUS", "ASYNCHRONOUS": "ASYNCHRONOUS",
                 "SEMISYNCHRONOUS": "SEMI_SYNCHRONOUS"}

    @classmethod
    def construct_server_entity_pb(cls, hostname, port, ssl_config_pb=None):
        return metis_pb2.ServerEntity(hostname=hostname, port=port, ssl_config=ssl_config_pb)

    @classmethod
    def construct_ssl_config_pb(cls, enable_ssl=False, config_pb=None):
        if config_pb is None:
            return metis_pb2.SSLConfig(enable_ssl=enable_ssl)
        return _wrap_oneof(metis_pb2.SSLConfig, config_pb, enable_ssl=enable_ssl)

    @classmethod
    def construct_ssl_config_files_pb(cls, public_certificate_file=None, private_key_file=None):
        return metis_pb2.SSLConfigFiles(public_certificate_file=public_certificate_file,
                                        private_key_file=private_key_file)

    @classmethod
    def construct_ssl_config_stream_pb(cls, public_certificate_stream=None, private_key_stream=None):
        return metis_pb2.SSLConfigStream(public_certificate_stream=public_certificate_stream,
                                         private_key_stream=private_key_stream)

    @classmethod
    def construct_he_scheme_config_pb(cls, enabled=False, crypto_context_file=None, public_key_file=None,
                                      private_key_file=None, empty_scheme_config_pb=None,
                                      ckks_scheme_config_pb=None):
        scheme = empty_scheme_config_pb if empty_scheme_config_pb is not None else ckks_scheme_config_pb
        base = dict(enabled=enabled, crypto_context_file=crypto_context_file, public_key_file=public_key_file,
                    private_key_file=private_key_file)
        if scheme is None:
            return metis_pb2.HESchemeConfig(**base)
        return _wrap_oneof(metis_pb2.HESchemeConfig, scheme, **base)

    @classmethod
    def construct_empty_scheme_config_pb(cls):
        return metis_pb2.EmptySchemeConfig()

    @classmethod
    def construct_ckks_scheme_config_pb(cls, batch_size, scaling_factor_bits):
        return metis_pb2.CKKSSchemeConfig(batch_size=batch_size, scaling_factor_bits=scaling_factor_bits)

    @classmethod
    def construct_dataset_spec_pb(cls, num_training_examples, num_validation_examples, num_test_examples,
                                  training_spec, validation_spec, test_spec,
                                  is_classification=False, is_regression=False):
        if not (is_classification or is_regression):
            raise RuntimeError("Need to specify whether incoming dataset spec is regression or classification.")
        kind = "classification" if is_classification else "regression"
        make = cls.construct_classification_dataset_spec_pb if is_classification \
            else cls.construct_regression_dataset_spec_pb
        specs = {f"{split}_{kind}_spec": make(spec) for split, spec in
                 (("training", training_spec), ("validation", validation_spec), ("test", test_spec))}
        return metis_pb2.DatasetSpec(num_training_examples=num_training_examples,
                                     num_validation_examples=num_validation_examples,
                                     num_test_examples=num_test_examples, **specs)

    @classmethod
    def construct_classification_dataset_spec_pb(cls, class_distribution_specs=None):
        return metis_pb2.DatasetSpec.ClassificationDatasetSpec(
            class_examples_num={int(k): int(v) for k, v in (class_distribution_specs or {}).items()})

    @classmethod
    def construct_regression_dataset_spec_pb(cls, regression_specs=None):
        r = regression_specs or {}
        keys = ("min", "max", "mean", "median", "mode", "stddev")
        return metis_pb2.DatasetSpec.RegressionDatasetSpec(**{k: float(r.get(k, 0.0)) for k in keys})

    @classmethod
    def construct_learning_task_pb(cls, num_local_updates, validation_dataset_pct, metrics=None):
        return metis_pb2.LearningTask(
            num_local_updates=num_local_updates,
            training_dataset_percentage_for_stratified_validation=validation_dataset_pct, metrics=metrics)

    @classmethod
    def construct_completed_learning_task_pb(cls, model_pb, task_execution_metadata_pb, aux_metadata):
        return metis_pb2.CompletedLearningTask(model=model_pb, execution_metadata=task_execution_metadata_pb,
                                               aux_metadata=aux_metadata)

    @classmethod
    def construct_task_execution_metadata_pb(cls, global_iteration, task_evaluation_pb, completed_epochs,
                                             completed_batches, batch_size, processing_ms_per_epoch,
                                             processing_ms_per_batch):
        return metis_pb2.TaskExecutionMetadata(
            global_iteration=global_iteration, task_evaluation=task_evaluation_pb,
            completed_epochs=completed_epochs, completed_batches=completed_batches, batch_size=batch_size,
            processing_ms_per_epoch=processing_ms_per_epoch, processing_ms_per_batch=processing_ms_per_batch)

    @classmethod
    def construct_task_evaluation_pb(cls, epoch_training_evaluations_pbs, epoch_validation_evaluations_pbs,
                                     epoch_test_evaluations_pbs):
        return metis_pb2.TaskEvaluation(training_evaluation=epoch_training_evaluations_pbs,
                                        validation_evaluation=epoch_validation_evaluations_pbs,
                                        test_evaluation=epoch_test_evaluations_pbs)

    @classmethod
    def construct_epoch_evaluation_pb(cls, epoch_id, model_evaluation_pb):
        return metis_pb2.EpochEvaluation(epoch_id=epoch_id, model_evaluation=model_evaluation_pb)

    @classmethod
    def construct_evaluation_metrics_pb(cls, metrics=None):
        if metrics is None:
            metrics = []
        elif not isinstance(metrics, (list, tuple)):
            metrics = [metrics]
        return metis_pb2.EvaluationMetrics(metric=[str(m) for m in metrics])

    @classmethod
    def construct_model_evaluation_pb(cls, metric_values=None):
        return metis_pb2.ModelEvaluation(metric_values={str(k): str(v) for k, v in (metric_values or {}).items()})

    @classmethod
    def construct_model_evaluations_pb(cls, training_evaluation_pb, validation_evaluation_pb, test_evaluation_pb):
        return metis_pb2.ModelEvaluations(training_evaluation=training_evaluation_pb,
                                          validation_evaluation=validation_evaluation_pb,
                                          test_evaluation=test_evaluation_pb)

    @classmethod
    def construct_hyperparameters_pb(cls, batch_size, optimizer_config_pb):
        return metis_pb2.Hyperparameters(batch_size=batch_size, optimizer=optimizer_config_pb)

    @classmethod
    def construct_controller_params_pb(cls, server_entity_pb, global_model_specs_pb, communication_specs_pb,
                                       model_store_config_pb, model_hyperparams_pb):
        return metis_pb2.ControllerParams(server_entity=server_entity_pb, global_model_specs=global_model_specs_pb,
                                          communication_specs=communication_specs_pb,
                                          model_store_config=model_store_config_pb,
                                          model_hyperparams=model_hyperparams_pb)

    @classmethod
    def construct_controller_modelhyperparams_pb(cls, batch_size, epochs, optimizer_pb, percent_validation):
        return metis_pb2.ControllerParams.ModelHyperparams(batch_size=batch_size, epochs=epochs,
                                                           optimizer=optimizer_pb,
                                                           percent_validation=percent_validation)

    @classmethod
    def construct_no_eviction_pb(cls):
        return metis_pb2.NoEviction()

    @classmethod
    def construct_lineage_length_eviction_pb(cls, lineage_length):
        return metis_pb2.LineageLengthEviction(lineage_length=lineage_length)

    @classmethod
    def construct_eviction_policy_pb(cls, policy_name, lineage_length):
        key = _norm(policy_name)
        if key == "NOEVICTION":
            return cls.construct_no_eviction_pb()
        if key == "LINEAGELENGTHEVICTION":
            return cls.construct_lineage_length_eviction_pb(lineage_length)
        raise RuntimeError(f"Not a supported eviction policy: {policy_name}")

    @classmethod
    def construct_model_store_specs_pb(cls, eviction_policy_pb):
        return _wrap_oneof(metis_pb2.ModelStoreSpecs, eviction_policy_pb)

    @classmethod
    def construct_model_store_config_pb(cls, name, eviction_policy, lineage_length=None, store_hostname=None,
                                        store_port=None):
        specs = cls.construct_model_store_specs_pb(cls.construct_eviction_policy_pb(eviction_policy, lineage_length))
        key = _norm(name)
        if key == "INMEMORY":
            return metis_pb2.ModelStoreConfig(in_memory_store=cls.construct_in_memory_store_pb(specs))
        if key == "REDIS":
            return metis_pb2.ModelStoreConfig(
                redis_db_store=cls.construct_redis_store_pb(specs, store_hostname, store_port))
        raise RuntimeError("Not a supported model store.")

    @classmethod
    def construct_in_memory_store_pb(cls, model_store_specs_pb):
        return metis_pb2.InMemoryStore(model_store_specs=model_store_specs_pb)

    @classmethod
    def construct_redis_store_pb(cls, model_store_specs_pb, hostname, port):
        return metis_pb2.RedisDBStore(model_store_specs=model_store_specs_pb,
                                      server_entity=cls.construct_server_entity_pb(hostname, port))

    @classmethod
    def construct_fed_avg_pb(cls):
        return metis_pb2.FedAvg()

    @classmethod
    def construct_fed_stride_pb(cls, stride_length):
        return metis_pb2.FedStride(stride_length=stride_length)

    @classmethod
    def construct_fed_rec_pb(cls):
        return metis_pb2.FedRec()

    @classmethod
    def construct_fed_async_pb(cls, alpha=0.5, staleness_fn="POLYNOMIAL", a=0.5, b=4.0):
        fn = staleness_fn if isinstance(staleness_fn, int) else \
            metis_pb2.FedAsync.StalenessFunction.Value(_norm(staleness_fn))
        return metis_pb2.FedAsync(alpha=alpha, staleness_fn=fn, a=a, b=b)

    @classmethod
    def construct_pwa_pb(cls, he_scheme_config_pb):
        return metis_pb2.PWA(he_scheme_config=he_scheme_config_pb)

    @classmethod
    def construct_aggregation_rule_specs_pb(cls, scaling_factor):
        try:
            value = metis_pb2.AggregationRuleSpecs.ScalingFactor.Value(cls._SCALING[_norm(scaling_factor)])
        except KeyError:
            raise RuntimeError("Unsupported scaling factor.") from None
        return metis_pb2.AggregationRuleSpecs(scaling_factor=value)

    @classmethod
    def construct_aggregation_rule_pb(cls, rule_name, scaling_factor, stride_length, he_scheme_config_pb,
                                      fed_async_kwargs=None):
        specs = cls.construct_aggregation_rule_specs_pb(scaling_factor)
        key = _norm(rule_name)
        if key == "FEDAVG":
            rule = cls.construct_fed_avg_pb()
        elif key == "FEDSTRIDE":
            rule = cls.construct_fed_stride_pb(stride_length)
        elif key == "FEDREC":
            rule = cls.construct_fed_rec_pb()
        elif key == "FEDASYNC":
            rule = cls.construct_fed_async_pb(**(fed_async_kwargs or {}))
        elif key == "PWA":
            rule = cls.construct_pwa_pb(he_scheme_config_pb=he_scheme_config_pb)
        else:
            raise RuntimeError("Unsupported rule name.")
        return _wrap_oneof(metis_pb2.AggregationRule, rule, aggregation_rule_specs=specs)

    @classmethod
    def construct_global_model_specs(cls, aggregation_rule_pb, learners_participation_ratio):
        return metis_pb2.GlobalModelSpecs(aggregation_rule=aggregation_rule_pb,
                                          learners_participation_ratio=learners_participation_ratio)

    @classmethod
    def construct_communication_specs_pb(cls, protocol, semi_sync_lambda=None, semi_sync_recompute_num_updates=None,
                                         task_deadline_secs=None):
        name = cls._PROTOCOL.get(_norm(protocol), "UNKNOWN")
        return metis_pb2.CommunicationSpecs(
            protocol=metis_pb2.CommunicationSpecs.Protocol.Value(name),
            protocol_specs=metis_pb2.ProtocolSpecs(
                semi_sync_lambda=semi_sync_lambda, semi_sync_recompute_num_updates=semi_sync_recompute_num_updates,
                task_deadline_secs=task_deadline_secs))


class ModelProtoMessages(object):

    class TensorSpecProto(object):
        """numpy / torch  <->  ``TensorSpec`` (reference :399-505, plus FLOAT16/BFLOAT16 and torch)."""

        @classmethod
        def _byte_order(cls, np_dtype):
            D = model_pb2.DType
            if np_dtype.itemsize == 1:
                return D.NA
            bo = np_dtype.byteorder
            if bo == "=":
                bo = "<" if sys.byteorder == "little" else ">"
            return D.LITTLE_ENDIAN_ORDER if bo == "<" else D.BIG_ENDIAN_ORDER

        @classmethod
        def numpy_array_to_proto_tensor_spec(cls, arr):
            if torch is not None and isinstance(arr, torch.Tensor):
                return cls.torch_tensor_to_proto_tensor_spec(arr)
            arr = np.asarray(arr)
            code = dtypes.numpy_to_wire(arr.dtype)
            fortran = bool(arr.flags.f_contiguous and not arr.flags.c_contiguous)
            flat = arr.flatten(order="F" if fortran else "C")
            return model_pb2.TensorSpec(
                length=flat.size, dimensions=list(arr.shape),
                type=model_pb2.DType(type=code, byte_order=cls._byte_order(arr.dtype), fortran_order=fortran),
                value=flat.tobytes())

        @classmethod
        def torch_tensor_to_proto_tensor_spec(cls, t):
            """Works for every wire dtype including bfloat16 (which numpy cannot represent)."""
            t = t.detach().contiguous().cpu()
            code = dtypes.torch_to_wire(t.dtype)
            raw = t.reshape(-1).view(torch.uint8).numpy().tobytes() if t.numel() else b""
            D = model_pb2.DType
            bo = D.NA if t.element_size() == 1 else (D.LITTLE_ENDIAN_ORDER if sys.byteorder == "little"
                                                     else D.BIG_ENDIAN_ORDER)
            return model_pb2.TensorSpec(length=t.numel(), dimensions=list(t.shape),
                                        type=D(type=code, byte_order=bo, fortran_order=False), value=raw)

        @classmethod
        def get_numpy_data_type_from_tensor_spec(cls, tensor_spec):
            D = model_pb2.DType
            dt = dtypes.wire_to_numpy(tensor_spec.type.type)
            if tensor_spec.type.byte_order == D.BIG_ENDIAN_ORDER:
                return dt.newbyteorder(">")
            if tensor_spec.type.byte_order == D.LITTLE_ENDIAN_ORDER:
                return dt.newbyteorder("<")
            return dt

        @classmethod
        def proto_tensor_spec_to_numpy_array(cls, tensor_spec):
            """bfloat16 tensors come back as their raw uint16 bits (numpy has no bf16)."""
            dt = cls.get_numpy_data_type_from_tensor_spec(tensor_spec)
            flat = np.frombuffer(tensor_spec.value, dtype=dt, count=tensor_spec.length)
            order = "F" if tensor_spec.type.fortran_order else "C"
            return flat.reshape(tuple(tensor_spec.dimensions), order=order)

        @classmethod
        def proto_tensor_spec_to_torch_tensor(cls, tensor_spec, device=None):
            arr = cls.proto_tensor_spec_to_numpy_array(tensor_spec)
            if arr.dtype.byteorder == ">" or (arr.dtype.byteorder == "=" and sys.byteorder == "big"):
                arr = arr.astype(arr.dtype.newbyteorder("<"))
            arr = np.ascontiguousarray(arr)
            target = dtypes.wire_to_torch(tensor_spec.type.type)
            if arr.size == 0:
                return torch.empty(tuple(tensor_spec.dimensions), dtype=target, device=device)
            t = torch.from_numpy(arr.view(np.uint8).reshape(-1).copy()).view(target).reshape(
                tuple(tensor_spec.dimensions))
            return t.to(device) if device is not None else t

        @classmethod
        def proto_tensor_spec_with_list_values_to_numpy_array_like(cls, tensor_spec, list_of_values):
            """Decrypted float64 values -> torch tensor of the variable's declared dtype and shape."""
            target = dtypes.wire_to_torch(tensor_spec.type.type)
            t = torch.as_tensor(np.asarray(list_of_values, dtype=np.float64))
            if not target.is_floating_point:
                t = t.round()
            return t.to(target).reshape(tuple(tensor_spec.dimensions))

        @classmethod
        def proto_tensor_spec_with_list_values_to_numpy_array(cls, tensor_spec, list_of_values):
            dt = cls.get_numpy_data_type_from_tensor_spec(tensor_spec)
            order = "F" if tensor_spec.type.fortran_order else "C"
            return np.array(list_of_values, dtype=dt).reshape(tuple(tensor_spec.dimensions), order=order)

    @classmethod
    def construct_tensor_pb(cls, nparray, ciphertext=None):
        if not (isinstance(nparray, np.ndarray) or (torch is not None and isinstance(nparray, torch.Tensor))):
            raise TypeError("Parameter {} must be of type {}.".format(type(nparray), np.ndarray))
        spec = cls.TensorSpecProto.numpy_array_to_proto_tensor_spec(nparray)
        if ciphertext is not None:  # the ciphertext replaces the payload; shape/dtype describe the plaintext
            spec.value = ciphertext
            return model_pb2.CiphertextTensor(tensor_spec=spec)
        return model_pb2.PlaintextTensor(tensor_spec=spec)

    @classmethod
    def construct_model_variable_pb(cls, name, trainable, tensor_pb):
        assert isinstance(name, str) and isinstance(trainable, (bool, np.bool_))
        return _wrap_oneof(model_pb2.Model.Variable, tensor_pb, name=name, trainable=bool(trainable))

    @classmethod
    def construct_model_pb_from_vars_pb(cls, variables_pb):
        assert all(isinstance(v, model_pb2.Model.Variable) for v in variables_pb)
        return model_pb2.Model(variables=list(variables_pb))

    @classmethod
    def construct_model_pb_from_np(cls, weights_values, weights_names, weights_trainable, he_scheme=None):
        variables = []
        for name, trainable, value in zip(weights_names, weights_trainable, weights_values):
            ciphertext = None
            if he_scheme is not None:
                flat = value.detach().float().cpu().numpy().ravel() if (torch is not None and isinstance(
                    value, torch.Tensor)) else np.asarray(value).ravel()
                ciphertext = he_scheme.encrypt(flat.astype(np.float64))
            variables.append(cls.construct_model_variable_pb(
                name=name, trainable=trainable, tensor_pb=cls.construct_tensor_pb(value, ciphertext=ciphertext)))
        return model_pb2.Model(variables=variables)

    @classmethod
    def construct_model_bytes_from_np(cls, weights_values, weights_names, weights_trainable, he_scheme=None) -> bytes:
        """The serialized ``Model`` of ``construct_model_pb_from_np`` without the protobuf objects in between: the core's
        wire writer copies every tensor once, straight out of the array's buffer (``_core.build_model``).  The reference
        builds ``TensorSpec`` -> ``PlaintextTensor`` -> ``Variable`` -> ``Model`` and serializes (:419-465, :543-564): a
        copy of every tensor per level."""
        from metisfl_b200 import _core
        D = model_pb2.DType
        variables = []
        for name, trainable, value in zip(weights_names, weights_trainable, weights_values):
            if torch is not None and isinstance(value, torch.Tensor):
                t = value.detach().contiguous().cpu()
                code, dims, length, fortran = dtypes.torch_to_wire(t.dtype), list(t.shape), t.numel(), False
                order = D.NA if t.element_size() == 1 else (D.LITTLE_ENDIAN_ORDER if sys.byteorder == "little"
                                                            else D.BIG_ENDIAN_ORDER)
                raw = t.reshape(-1).view(torch.uint8).numpy() if t.numel() else b""
                plain = t
            else:
                arr = np.asarray(value)
                code, dims, length = dtypes.numpy_to_wire(arr.dtype), list(arr.shape), arr.size
                fortran = bool(arr.flags.f_contiguous and not arr.flags.c_contiguous)
                order = cls.TensorSpecProto._byte_order(arr.dtype)
                flat = np.ascontiguousarray(arr.reshape(-1, order="F" if fortran else "C"))
                raw = flat.view(np.uint8) if flat.size else b""
                plain = arr
            entry = {"name": name, "trainable": bool(trainable), "length": int(length), "dimensions": dims,
                     "dtype": int(code), "byte_order": int(order), "fortran_order": fortran, "value": raw}
            if he_scheme is not None:   # the ciphertext replaces the payload; shape / dtype describe the plaintext
                flat64 = plain.detach().float().cpu().numpy().ravel() if (torch is not None and isinstance(
                    plain, torch.Tensor)) else np.asarray(plain).ravel()
                entry["value"] = he_scheme.encrypt(flat64.astype(np.float64))
                entry["ciphertext"] = True
            variables.append(entry)
        return _core.build_model(variables)

    @classmethod
    def construct_federated_model_pb(cls, num_contributors, model_pb):
        assert isinstance(model_pb, model_pb2.Model)
        return model_pb2.FederatedModel(num_contributors=num_contributors, model=model_pb)

    @classmethod
    def construct_vanilla_sgd_optimizer_pb(cls, learning_rate, l1_reg=0.0, l2_reg=0.0):
        return model_pb2.VanillaSGD(learning_rate=learning_rate, L1_reg=l1_reg, L2_reg=l2_reg)

    @classmethod
    def construct_momentum_sgd_optimizer_pb(cls, learning_rate, momentum_factor=0.0):
        return model_pb2.MomentumSGD(learning_rate=learning_rate, momentum_factor=momentum_factor)

    @classmethod
    def construct_fed_prox_optimizer_pb(cls, learning_rate, proximal_term=0.0):
        return model_pb2.FedProx(learning_rate=learning_rate, proximal_term=proximal_term)

    @classmethod
    def construct_adam_optimizer_pb(cls, learning_rate, beta_1=0.0, beta_2=0.0, epsilon=0.0):
        return model_pb2.Adam(learning_rate=learning_rate, beta_1=beta_1, beta_2=beta_2, epsilon=epsilon)

    @classmethod
    def construct_adam_optimizer_with_weight_decay_pb(cls, learning_rate, weight_decay):
        return model_pb2.AdamWeightDecay(learning_rate=learning_rate, weight_decay=weight_decay)

    _OPTIMIZERS = {
        "VanillaSGD": "construct_vanilla_sgd_optimizer_pb", "MomentumSGD": "construct_momentum_sgd_optimizer_pb",
        "FedProx": "construct_fed_prox_optimizer_pb", "Adam": "construct_adam_optimizer_pb",
        "AdamWeightDecay": "construct_adam_optimizer_with_weight_decay_pb",
        "AdamW": "construct_adam_optimizer_with_weight_decay_pb",
    }

    @classmethod
    def construct_optimizer_config_pb_from_kwargs(cls, optimizer_pb_kwargs):
        kwargs = dict(optimizer_pb_kwargs)
        name = kwargs.pop("name", None)
        if name not in cls._OPTIMIZERS:
            raise RuntimeError("Optimizer kwargs refer to a non-supported optimizer.")
        return cls.construct_optimizer_config_pb(getattr(cls, cls._OPTIMIZERS[name])(**kwargs))

    @classmethod
    def construct_optimizer_config_pb(cls, optimizer_pb):
        return _wrap_oneof(model_pb2.OptimizerConfig, optimizer_pb)


class ServiceCommonProtoMessages(object):

    @classmethod
    def construct_ack_pb(cls, status, google_timestamp, message=None):
        return service_common_pb2.Ack(status=status, timestamp=google_timestamp, message=message)

    @classmethod
    def construct_get_services_health_status_request_pb(cls):
        return service_common_pb2.GetServicesHealthStatusRequest()

    @classmethod
    def construct_get_services_health_status_response_pb(cls, services_status):
        return service_common_pb2.GetServicesHealthStatusResponse(services_status=services_status)

    @classmethod
    def construct_shutdown_request_pb(cls):
        return service_common_pb2.ShutDownRequest()

    @classmethod
    def construct_shutdown_response_pb(cls, ack_pb):
        return service_common_pb2.ShutDownResponse(ack=ack_pb)
