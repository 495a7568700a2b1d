"""NetworkFactory mirror: the shape description of the policy-value network
(NetworkFactory.py:11-19 constructor arguments; the forward graph of :22-183 is the CUDA kernel
pair in csrc/net.cu / csrc/net_tc.cuh; the loss / optimizer half (:185-245) is csrc/train.cu)."""


class NetworkFactory(object):
    def __init__(self, networkConfig, policyShape):
        self.networkConfig = networkConfig
        self.policyShape = policyShape
        self.blocks = int(networkConfig.get('blocks', 4))                       # parameters_template.yaml:12-16
        self.filters = int(networkConfig.get('filters', 16))
        self.dense = int(networkConfig.get('eval', {}).get('dense', 16))
        noise = networkConfig.get('policy', {}).get('dirichlet', {})
        self.epsilon = float(noise.get('epsilon', 0.0))
        self.alpha = float(noise.get('alpha', 0.2))
        training = networkConfig.get('training', {}) or {}
        self.optimizer = str(training.get('optimizer', 'sgd'))                    # NetworkFactory.py:229-242: adam | momentum | anything else = SGD
        self.momentum = float(training.get('momentum', 0.9) or 0.0)
        self.hasTeacher = bool(networkConfig.get('hasTeacher'))                   # NetworkFactory.py:18,206
