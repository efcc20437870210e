"""``run_miniapp`` of the reference's densityrw/rw_miniapp_fugu.py:22-60 on a B200.

The reference lays the DensityRW brick on a Fugu scaffold, runs the spiking back end and returns
the spike table of the read-out neurons (pandas: ``time``, ``neuron_number`` + a ``node`` column of
location strings) from which ``fluence_mini_app.run_fugu`` (:72-107) recovers walkers per (node,
model step): a new model step starts wherever consecutive spike times differ by more than 1.  This
version walks on the GPU and writes the same table: one row per walker per model step, model step s
at tick 2*s, so the reference's own post-processing reproduces the walk's counts."""
import argparse
import logging

import numpy as np

from .DensityRW import DensityRW, load_transitions, transitions_to_csr

rw_logger = logging.getLogger("NRW")


def init_walker_tuple(s):
    try:
        node_number, walkers = map(int, s.split(','))
        return (node_number, walkers)
    except Exception:
        raise argparse.ArgumentTypeError("Must be (node_number, walkers) pair")


def counts_to_spike_table(counts, include_rows=None):
    """counts[n_nodes, n_steps] -> DataFrame(time, neuron_number, node), sorted by time."""
    import pandas as pd
    c = np.asarray(counts).astype(np.int64)
    if include_rows is not None:
        c = c.copy()
        c[~np.asarray(include_rows, dtype=bool)] = 0
    nodes, steps = np.nonzero(c)
    order = np.lexsort((nodes, steps))
    nodes, steps = nodes[order], steps[order]
    reps = c[nodes, steps]
    node_col = np.repeat(nodes, reps)
    time_col = np.repeat(2 * steps, reps).astype(np.float64)
    return pd.DataFrame({"time": time_col, "neuron_number": node_col.astype(np.float64),
                         "node": ["(%d,)" % k for k in node_col]})


def run_miniapp(transitions_file=None, neural_timesteps=100, initial_walkers=[],
                sink_connections=False, backend='snn', rng=None, sampler="alias"):
    """rw_miniapp_fugu.py:22-60.  ``backend`` 'snn' (the reference's CPU spiking simulator) and
    'b200' both run the GPU walk here; 'loihi' needs NxSDK and is refused.  ``neural_timesteps``
    counts model steps.  With ``sink_connections=False`` probability-1 self-loops are removed (a sink
    absorbs) and, as in the reference, walkers inside a sink are not reported."""
    if backend not in ('snn', 'b200'):
        raise ValueError("Unsupported backend type.")
    transitions = load_transitions(transitions_file, remove_sink_connections=not sink_connections,
                                   verbose=0)
    walk = DensityRW(neural_timesteps, transitions,
                     init_walkers={(t[0],): t[1] for t in initial_walkers}, name='DensityRW')
    counts = walk.run(rng=rng, sampler=sampler)
    csr = walk.csr
    tracked = np.ones(csr.n_rows, dtype=bool)
    if not sink_connections:
        tracked = csr.row_lengths() > 0
    return counts_to_spike_table(counts, include_rows=tracked)


if __name__ == '__main__':
    parser = argparse.ArgumentParser()
    parser.add_argument('--transitions_file', required=True)
    parser.add_argument('--neural_timesteps', type=int, default=100)
    parser.add_argument('--remove_sink_connections', default=False, action='store_false', dest='sink_connections')
    parser.add_argument('--leave_sink_connections', action='store_true', dest='sink_connections')
    parser.add_argument('--initial_walkers', type=init_walker_tuple, nargs='*')
    parser.add_argument('--backend', choices=['snn', 'b200'], default='snn')
    args = vars(parser.parse_args())
    rw_logger.info("Using parameters: {}".format(args))
    result = run_miniapp(**args)
    rw_logger.info(result)
