"""TEST INFRASTRUCTURE -- tests/golden/ga3c_probs.npz: the GA3C-CADRL network evaluated on the reference's SHIPPED
weights, by a numpy evaluation of the TensorFlow graph written from `GA3C_CADRL/network.py` itself and from the
documented semantics of the TF 1.15 ops it calls -- independently of oracle/casim_oracle_ga3c.c (no shared code, a
different language, float64 AND float32 arithmetic), so that the C oracle and the CUDA kernels are checked against a
second derivation instead of against themselves.

    python oracle/gen_golden_ga3c_probs.py          (build container; needs /root/reference for the checkpoint)

What is evaluated, op by op (gym_collision_avoidance/envs/policies/GA3C_CADRL/network.py):
  :74-77   x_normalized = (x - NN_INPUT_AVG_VECTOR) / NN_INPUT_STD_VECTOR              (constants :125-148)
  :85      num_other_agents = x[:, 0]                 (un-normalised; dynamic_rnn casts sequence_length to int32)
  :86-88   host_agent_vec = x_normalized[:, 1:5]; other_agent_seq = reshape(x_normalized[:, 5:], [-1, 10, 7])
  :89      tf.nn.dynamic_rnn(tf.contrib.rnn.LSTMCell(64), seq, sequence_length=num_other_agents): zero initial state;
           for t >= sequence_length the state is copied through; tf.contrib.rnn.LSTMCell.call (rnn_cell_impl.py):
             lstm_matrix = matmul(concat([inputs, m_prev], 1), kernel) + bias;  i, j, f, o = split(lstm_matrix, 4, 1)
             c = sigmoid(f + forget_bias(1.0)) * c_prev + sigmoid(i) * tanh(j);  m = sigmoid(o) * tanh(c)
  :90      rnn_output = rnn_state.h  (= m of the last valid step)
  :91-92   layer1 = relu(dense(concat([host_agent_vec, rnn_output], 1)));  :94 layer2 = relu(dense(layer1))
  :45-51   fullyconnected1 = relu(dense(flatten(layer2))); logits_p = dense(fc1);
           softmax_p = (softmax(logits_p) + MIN_POLICY(1e-4)) / (1 + MIN_POLICY * num_actions)
  GA3CCADRLPolicy.py:38-39   action_index = argmax(predictions)

Inputs: the state vectors of tests/golden/ga3c_state.npz -- produced by the UNMODIFIED reference function
agents_to_ga3c_cadrl_state (oracle/gen_golden_ga3c.py) -- cast to float32 like the tf.float32 placeholder (:43).
Weights: read by TF variable NAME from the reference's checkpoint GA3C_CADRL/checkpoints/IROS18/network_01900000
(the default of GA3CCADRLPolicy.initialize_network, GA3CCADRLPolicy.py:21-32).

What this does NOT pin: TensorFlow's own float32 rounding (TF 1.15 is not installable here).  The fixture stores the
float64 result ("the graph's exact value") and the float32 result of the same op order; they differ by < 2e-6, which
bounds what any float32 implementation, TF's included, can differ by -- hence the tolerances in the tests.
"""
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
GOLDEN = os.path.join(_ROOT, "tests", "golden")

AVG = np.array([0.0] + [0.0, 0.0, 1.0, 0.5] + [0.0, 0.0, 0.0, 0.0, 0.5, 0.0, 1.0] * 10)     # network.py:125-148
STD = np.array([1.0] + [5.0, 3.14, 1.0, 1.0] + [5.0, 5.0, 1.0, 1.0, 1.0, 5.0, 1.0] * 10)


def tf_graph(x, v, dtype):
    """x [B,75]; v: dict TF variable name -> array.  Every op in `dtype`."""
    f = lambda a: np.asarray(a, dtype=dtype)           # noqa: E731
    sig = lambda a: f(1) / (f(1) + np.exp(-a))         # noqa: E731
    x = f(x)
    B = x.shape[0]
    xn = (x - f(AVG)) / f(STD)
    seq_len = x[:, 0].astype(np.int32)
    host = xn[:, 1:5]
    seq = xn[:, 5:].reshape(B, 10, 7)
    K, b = f(v["rnn/lstm_cell/kernel"]), f(v["rnn/lstm_cell/bias"])
    c = np.zeros((B, 64), dtype)
    m = np.zeros((B, 64), dtype)
    for t in range(10):
        z = np.concatenate([seq[:, t], m], axis=1) @ K + b
        i, j, fg, o = z[:, 0:64], z[:, 64:128], z[:, 128:192], z[:, 192:256]
        c_new = sig(fg + f(1.0)) * c + sig(i) * np.tanh(j)
        m_new = sig(o) * np.tanh(c_new)
        live = (t < seq_len)[:, None]
        c = np.where(live, c_new, c)
        m = np.where(live, m_new, m)
    a = np.concatenate([host, m], axis=1)
    for name in ("layer1", "layer2", "fullyconnected1"):
        a = np.maximum(a @ f(v[name + "/kernel"]) + f(v[name + "/bias"]), f(0))
    logits = a @ f(v["logits_p/kernel"]) + f(v["logits_p/bias"])
    e = np.exp(logits - logits.max(axis=1, keepdims=True))
    sm = e / e.sum(axis=1, keepdims=True)
    return (sm + f(1e-4)) / (f(1.0) + f(1e-4) * f(11))


def main():
    sys.path.insert(0, _ROOT)
    sys.path.insert(0, os.path.join(_ROOT, "gym-exploration-2d_b200"))
    from oracle import ref_harness as rh
    from casim.ga3c_checkpoint import read_checkpoint
    prefix = os.path.join(rh.REFERENCE_ROOT, "gym_collision_avoidance", "envs", "policies", "GA3C_CADRL", "checkpoints",
                          "IROS18", "network_01900000")
    t = read_checkpoint(prefix)
    v = {k.split(":")[0]: np.asarray(a) for k, a in t.items()}
    z = np.load(os.path.join(GOLDEN, "ga3c_state.npz"))
    xs = []
    for k in range(z["nn_input"].shape[0]):
        n = int(z["n_agents"][k])
        xs.append(z["nn_input"][k, :n])
    x = np.concatenate(xs).astype(np.float32)                   # the tf.float32 placeholder
    p64 = tf_graph(x, v, np.float64)
    p32 = tf_graph(x, v, np.float32)
    srt = np.sort(p64, axis=1)
    np.savez_compressed(os.path.join(GOLDEN, "ga3c_probs.npz"), x=x, probs64=p64, probs32=p32.astype(np.float32),
                        action=p64.argmax(1).astype(np.int32), margin=srt[:, -1] - srt[:, -2])
    print("wrote ga3c_probs.npz: %d state vectors; max |p32 - p64| = %.3g; argmax(p32) != argmax(p64): %d; "
          "min top-2 margin %.3g" % (len(x), np.abs(p32 - p64).max(), int((p32.argmax(1) != p64.argmax(1)).sum()),
                                      (srt[:, -1] - srt[:, -2]).min()))


if __name__ == "__main__":
    main()
