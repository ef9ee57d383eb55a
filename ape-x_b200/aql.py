"""AQL candidate-action Q evaluation on the B200 tensor cores.

Mirror of the per-candidate part of `Q_Network.act` (model.py:294-335): given the state
feature f = q_feature(state) [B, 64] (a per-state MLP that stays PyTorch) and the K candidate
actions per state `a_mu`, returns q_values [B, K] and argmax_k — what `AQL.forward`
(model.py:193-196) and `utils.compute_loss_AQL` (utils.py:47-49) consume — without building
the [B, K, 128] concat tensor.  Weights are handed over once per optimiser step
(`set_weights`); NoisyLinear layers contribute their effective weights mu + sigma * epsilon in
training mode, mu otherwise (model.py:132-143).
"""
import torch

import per_native as _n


class CandidateQ(object):
    def __init__(self, continuous, action_dim=1, device=None):
        self.continuous = bool(continuous)
        self.action_dim = int(action_dim) if continuous else 1
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        h = _n.C.c_void_p()
        _n.check(_n.lib.aql_create(_n.C.byref(h), self.device.index, int(self.continuous), self.action_dim))
        self._h = h
        self._wadv1 = self._badv1 = None

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and _n is not None:
            _n.lib.aql_destroy(h)
            self._h = None

    @staticmethod
    def noisy_effective(layer):
        """Effective (weight, bias) of a NoisyLinear (model.py:132-143)."""
        if layer.training:
            return (layer.weight_mu + layer.weight_sigma * layer.weight_epsilon.to(layer.weight_mu.device),
                    layer.bias_mu + layer.bias_sigma * layer.bias_epsilon.to(layer.bias_mu.device))
        return layer.weight_mu, layer.bias_mu

    def set_weights(self, wa1, ba1, wa2, ba2, wadv1, badv1, wadv2, badv2):
        """float32 CUDA tensors in torch's [out, in] layout.  continuous: wa1 [128, A], ba1 [128],
        wa2 [64, 128], ba2 [64]; discrete: wa1 [64, 1], ba1 [64], wa2 = ba2 = None.  wadv1 [64, 128],
        badv1 [64], wadv2 [1, 64], badv2 [1] are the effective advantage weights."""
        prep = lambda t: None if t is None else t.detach().to(self.device, torch.float32).contiguous()
        ws = [prep(t) for t in (wa1, ba1, wa2, ba2, wadv1, badv1, wadv2, badv2)]
        _n.check(_n.lib.aql_set_weights(self._h, *[_n.ptr(t) for t in ws], _n.stream_ptr()))
        self._wadv1, self._badv1 = ws[4], ws[5]
        self._keep = ws

    def set_weights_from(self, q_network):
        """Reads the layers of a reference `Q_Network` (model.py:209-281)."""
        ao = q_network.action_out
        w1, b1 = self.noisy_effective(q_network.advantage1)
        w2, b2 = self.noisy_effective(q_network.advantage2)
        if self.continuous:
            self.set_weights(ao[0].weight, ao[0].bias, ao[2].weight, ao[2].bias, w1, b1, w2, b2)
        else:
            self.set_weights(ao[0].weight, ao[0].bias, None, None, w1, b1, w2, b2)

    def forward(self, f, a_mu, want_argmax=True):
        """f float32 [B, 64]; a_mu float32 [B, K, A] (continuous) or int64 [B, K] (discrete)."""
        B = int(f.shape[0])
        if self.continuous:
            a_mu = a_mu.reshape(B, -1, self.action_dim).to(torch.float32).contiguous()
        else:
            a_mu = a_mu.reshape(B, -1).to(torch.int64).contiguous()
        K = int(a_mu.shape[1])
        f = f.to(torch.float32).contiguous()
        q = torch.empty(B, K, dtype=torch.float32, device=self.device)
        am = torch.empty(B, dtype=torch.int64, device=self.device) if want_argmax else None
        _n.check(_n.lib.aql_candidate_q(self._h, _n.ptr(self._wadv1), _n.ptr(self._badv1), _n.ptr(f), _n.ptr(a_mu), B, K,
                                        _n.ptr(q), _n.ptr(am), _n.stream_ptr()))
        return q, am
