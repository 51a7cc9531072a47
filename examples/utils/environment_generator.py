"""Manufacture an N-learner localhost environment from a one-learner template (reference
``examples/utils/environment_generator.py:8-37``): consecutive ports, round-robin GPU assignment."""
import copy
import itertools

from metisfl_b200.utils.fedenv_parser import FederationEnvironment


class EnvGen(object):

    def __init__(self, template_filepath=None):
        self.template_filepath = template_filepath

    def generate_localhost(self, federation_rounds=10, learners_num=10, gpu_devices=(-1,), gpu_assignment="round_robin"):
        if gpu_assignment != "round_robin":
            raise RuntimeError("Only Round-Robin assignment is currently supported.")
        env = FederationEnvironment(self.template_filepath)
        env.termination_signals.federation_rounds = federation_rounds
        template = env.learners.learners[0]
        devices = itertools.cycle(gpu_devices)
        env.learners.learners = []
        for k in range(learners_num):
            learner = copy.deepcopy(template)
            learner.learner_id = "localhost:{}".format(k)
            d = int(next(devices))
            learner.cuda_devices = [d] if d >= 0 else []
            learner.grpc_servicer.port = int(template.grpc_servicer.port + k)
            env.learners.learners.append(learner)
        return env
