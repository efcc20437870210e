"""Two-state flux miniapp -- drop-in for the reference's ``simple_flux_miniapp.py`` on a B200.

States: 0 = positive direction, 1 = negative direction, 2 = absorbed (never stepped).  Same entry
points, command line (README.md:150-212 of the reference) and output files; ``--run_mode pc`` runs
on the GPU through ``nrw``.  Reference lines are cited as simple_flux_miniapp.py:<line>.
"""
from __future__ import absolute_import

import argparse
import logging
import os
import random
import signal
import sys
import time
from datetime import datetime

os.environ['SLURM'] = "1"

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import utils.logging_utils as utils       # noqa: E402
from nrw import dist as ndist, dropin, mesh   # noqa: E402
from nrw.engine import get_engine          # noqa: E402

rw_logger = logging.getLogger("NRW")


def signal_handler(signum, frame):
    rw_logger.info('Signal handler called with signal {}'.format(signum))
    os.killpg(os.getpid(), signum)
    sys.exit(0)


def run_pc(matrix_filename=None, neuraltime=None, initial_walkers=None, meshSize=None,
           runtime_args={}):
    """simple_flux_miniapp.py:174-203 on the GPU: float64 ``tempdata[2, neuraltime]`` (walkers in the
    + and - states per step; absorbed walkers vanish, Q9).  The `<=` threshold chain of :192-201 is
    searchsorted(left) on the row's 3-entry prefix sums, which is what the kernels evaluate."""
    csr = mesh.load_matrix(matrix_filename, n_live=2)
    return dropin.pc_tempdata(csr, neuraltime, initial_walkers, 2, rng=runtime_args.get('rng'),
                              sampler=runtime_args.get('sampler', 'alias'))


# The 'fugu' and 'loihi' entries keep the reference's adapters (fluence_mini_app.py:43-192) but sit
# on this build's densityrw / loihirw surfaces, i.e. the walk runs on the GPU and is written out as
# the spike table / spike probes those adapters decode.  (The command line still refuses
# ``-r loihi`` without NxSDK, as the reference does, :527-533.)
run_fugu = dropin.fugu_adapter
run_loihi = dropin.loihi_adapter
run_spinnaker = dropin.unavailable_backend('spinnaker', 'sPyNNaker and a SpiNNaker board')

run_functions = {'fugu': run_fugu, 'loihi': run_loihi, 'spinnaker': run_spinnaker, 'pc': run_pc}


def _counts_pc_fused(csr, M, nt, n_cols, rng, sampler):
    """Occupancy counts [start state, state, t < n_cols] for both start states in one pass; state 2 = walkers
    absorbed AT step t (the sink row of the 3-state chain, simple_flux_miniapp.py:197-201)."""
    eng, tabs = dropin.device_tables(csr)
    mode = dropin.resolve_rng(rng, 2 * int(M))
    td = np.zeros((2, 3, n_cols))
    if mode == "replay":
        # the reference walks until every walker is absorbed or nt is reached (:189); so does replay,
        # which keeps the global stream position identical
        rep = eng.replay(tabs, [0, 1], [M, M], nt)
        off = rep.ev_offsets.cpu().numpy()
        t = rep.ev_t[:rep.n_events].cpu().numpy()
        k = rep.ev_node[:rep.n_events].cpu().numpy()
        c = rep.ev_count[:rep.n_events].cpu().numpy()
        for s in (0, 1):
            sl = slice(int(off[s]), int(off[s + 1]))
            keep = (k[sl] < 3) & (t[sl] < n_cols)
            td[s, k[sl][keep], t[sl][keep]] = c[sl][keep]
        return td, rep.draws_used, mode
    seed = ndist.broadcast_int(dropin.philox_seed_from_global_stream())
    nt_k = max(min(int(nt), n_cols), 1)       # steps past the reported window cannot change it
    # every rank walks its slice of both start states' walkers; one all-reduce of the counts
    rank, world = ndist.world()
    lo, hi = ndist.shard_range(int(M), rank, world)
    res = eng.walk(tabs, [0, 1], [hi - lo, hi - lo], nt_k, seed, sampler=sampler, count_tn=True,
                   t_cap=nt_k, count_per_start=True, start_id_base=[lo, int(M) + lo])
    ndist.allreduce_sum_([res.count_tn, res.counters])
    cnt = res.count_tn.cpu().numpy()          # [2, t_cap, 3]
    td[:, :, :nt_k] = np.transpose(cnt[:, :, :3], (0, 2, 1))
    td[0, 0, 0] = M
    td[1, 1, 0] = M
    return td, res.counters_dict()["steps"], mode


def run(args):
    """simple_flux_miniapp.py:207-359."""
    rmode = args['run_mode']
    M, dt = args['num_walks'], args['timestep']
    initp, initn = args['init_pos'], args['init_neg']
    SigS, SigA = args['SigS'], args['SigA']
    max_time, neuraltime = args['max_time'], args['neural_timesteps']
    matrix_filename, results_filename = args['matrix_filename'], args['results_filename']

    try:
        myQ = mesh.build_simple_flux_dense(dt, SigS, SigA)                    # :230-285
    except ValueError:
        rw_logger.error("The probability of more than one scattering event occurring in a single "
                        "time step is greater than 0.01.")
        rw_logger.error("Alter your choice of SigA, SigS, or dt so that "
                        "1-(SigA+SigS)*dt*exp(-(SigA+SigS)*dt)-exp(-(SigA+SigS)*dt)<=0.01.")
        raise
    meshSize = 2
    rank, world = ndist.init_from_env()        # under torchrun: one process per GPU, rank 0 writes the files
    if rank == 0 or rmode != 'pc':
        np.savetxt(matrix_filename, myQ, delimiter=",")                       # :290
    rw_logger.info("Transition matrix construction complete")
    rw_logger.info("    matrix saved to {}".format(matrix_filename))

    if rmode == 'pc':
        maxT = min(neuraltime, int(max_time / dt)) + 2                        # :296-297
    else:
        maxT = int(max_time / dt) + 2
    rwSol = np.zeros((2, maxT))
    rw_logger.info("Beginning simulation")
    rw_logger.info("    using {} on {}".format(rmode, datetime.now()))
    if rmode == 'pc' and run_functions['pc'] is run_pc:
        t0 = time.time()
        csr = mesh.csr_from_dense(myQ, n_live=2)
        maxT = max(min(neuraltime - 2, maxT), 0)                              # :324
        td, steps, mode = _counts_pc_fused(csr, M, neuraltime, max(maxT, 1), args.get('rng'),
                                           args.get('sampler', 'alias'))
        for i in range(meshSize):
            rwSol[i, :maxT] = (initp * td[i, 0, :maxT] + initn * td[i, 1, :maxT]) / M   # :327
        if args['use_sinks']:
            # the reference's pc path ignores --use_sinks (Q14); here it ALSO returns the walkers absorbed at every
            # step -- next to rwSol, never inside it: sinks[i, t] for start state i, kept in
            # ``last_sink_counts`` and written to <results stem>_sinks.csv
            global last_sink_counts
            last_sink_counts = td[:, 2, :maxT].copy()
            if rank == 0:
                np.savetxt(os.path.splitext(results_filename)[0] + "_sinks.csv", last_sink_counts, delimiter=",")
        el = time.time() - t0
        rw_logger.info('    2 mesh points, {} walker-steps, rng={} in {:.4f} s'.format(steps, mode, el))
    else:
        run_function = run_functions[rmode]
        for i in range(0, meshSize):
            t0 = time.time()
            tempdata = run_function(matrix_filename=matrix_filename, neuraltime=neuraltime,
                                    initial_walkers=[[i, M]], meshSize=meshSize,
                                    runtime_args={'fugu_backend': args['fugu_backend'],
                                                  'use_sinks': args['use_sinks'],
                                                  'rand_seed': args['rand_seed'],
                                                  'rng': args.get('rng'),
                                                  'sampler': args.get('sampler', 'alias')})
            rw_logger.info('    mesh point: {}'.format(i))
            rw_logger.debug('        elapsed time: {}'.format(time.time() - t0))
            maxT = min((tempdata.shape[1] - 2), maxT)
            rwSol[i, :maxT] = (initp * tempdata[0, :maxT] + initn * tempdata[1, :maxT]) / M
    rw_logger.info("Simulation using {} completed on {}".format(rmode, datetime.now()))

    if rank == 0:
        if args['plot_results']:
            plot_results(rwSol, maxT, args)
        mesh.save_array_csv(results_filename, rwSol)                          # :358 (np.savetxt's bytes)
        rw_logger.info("Solution saved to {}".format(results_filename))
    return rwSol


last_sink_counts = None     # float64 [2, maxT]: walkers absorbed at each step per start state (--use_sinks, pc mode)


def true_solution(args, maxT):
    """Closed form of the two-state problem (simple_flux_miniapp.py:338-343)."""
    dt, ip, in_ = args['timestep'], args['init_pos'], args['init_neg']
    SigS, SigA = args['SigS'], args['SigA']
    t = np.arange(maxT) * dt
    e1, e2 = np.exp(-SigA * t), np.exp(-(SigA + SigS) * t)
    return t, np.stack([(ip / 2) * (e1 + e2) + (in_ / 2) * (e1 - e2),
                        (ip / 2) * (e1 - e2) + (in_ / 2) * (e1 + e2)])


def plot_results(rwSol, maxT, args):
    try:
        import matplotlib
        matplotlib.use('Agg')
        import matplotlib.pyplot as plt
    except ImportError:
        rw_logger.warning("matplotlib is not installed: skipping the plot")
        return
    tarray, trueSol = true_solution(args, maxT)
    rw_logger.info("Plotting solution")
    f1 = plt.figure(1)
    plt.plot(tarray, trueSol[0, :], color='k', label='Analytic', linestyle='dashed')
    plt.plot(tarray, trueSol[1, :], color='k', linestyle='dashed')
    plt.plot(tarray, rwSol[0, 0:maxT], color='r', label='Positive')
    plt.plot(tarray, rwSol[1, 0:maxT], color='b', label='Negative')
    plt.xlabel("Time")
    plt.ylabel("Solution")
    plt.legend()
    figname = args['results_filename'].split(sep='.')[0] + '_plot.png'
    f1.savefig(figname)
    plt.close(f1)
    rw_logger.info("Plot saved to {}".format(figname))


_RUN_MODE_HELP = ('The mode specifies how the random walks are ran. Must be one of fugu, loihi, '
                  'spinnaker, or pc. The default is fugu. Specifying pc gives a python implementation.')
CLI = [
    (('-r', '--run_mode'), dict(default='fugu', choices=['fugu', 'loihi', 'spinnaker', 'pc'], help=_RUN_MODE_HELP)),
    (('-M', '--num_walks'), dict(type=int, default=500, help='The number of walkers to start on each location in the mesh.  The default is 500.')),
    (('-dt', '--timestep'), dict(type=float, default=0.005, help='The time step of the simulation. The default is 0.005.')),
    (('-ip', '--init_pos'), dict(type=float, default=5.0, help='The initial value for the positive direction solution. The default is 5.')),
    (('-in', '--init_neg'), dict(type=float, default=3.0, help='The initial value for the negative direction solution. The default is 3.')),
    (('-ss', '--SigS'), dict(type=float, default=8.0, help='The scattering cross-section of the particle. This is the rate at which scattering events occur. The default is 8.0.')),
    (('-sa', '--SigA'), dict(type=float, default=2.0, help='The absorption cross-section.  This is the rate at which particles are absorbed. The default is 2.0.')),
    (('-mt', '--max_time'), dict(type=float, default=5.0, help='The absolute max model time of simulation. This is the largest value of time that could be displayed in the solution. The actual largest model time of simulation may be smaller than this value if the neural timestep is too small for Loihi or Fugu->Loihi simulations. For conventional (pc) simulations, the actual time may be smaller if dt*nt<mt.')),
    (('-nt', '--neural_timesteps'), dict(type=int, default=500000, help='The number of neural timesteps the simulation is run for. This should be MUCH greater than the number of model time steps desired.  The default is 500000. For pc, this is the number of model timesteps.')),
    (('--debug',), dict(default=logging.WARNING, action="store_const", dest="loglevel", const=logging.DEBUG, help="Print logging DEBUG statements.")),
    (('--verbose',), dict(action="store_const", dest="loglevel", const=logging.INFO, help="Print logging INFO statements.")),
    (('--matrix_filename',), dict(default='curmat.csv', help='Filename for temporary storage of transition matrix.')),
    (('--results_filename',), dict(default='rwSol.csv', help='Filename to record the results.')),
    (('--fugu_backend',), dict(default='snn', choices=['snn'], help='Runs a spiking simulator using Fugu.')),
    (('--plot_results',), dict(default=False, action='store_true', help='Plots the results after finishing.')),
    (('--log_file',), dict(default='~/.miniapps/simple_flux_miniapp.log', help='Location of the log file.')),
    (('--rand_seed',), dict(type=int, default=-1, help="Sets the seed for the random number generator. Otherwise uses a time-based 'random' RNG seed.")),
    (('--use_sinks',), dict(default=False, action='store_true', help='Include sinks in a fugu simulation; default is False.')),
    # additive
    (('--rng',), dict(default='auto', choices=list(dropin.RNG_MODES), help="pc mode RNG: 'replay' = numpy legacy stream (bit-identical to the reference), 'philox' = on-GPU; 'auto' replays up to ~4M walkers.")),
    (('--sampler',), dict(default='alias', choices=list(dropin.SAMPLERS), help='philox mode transition sampler.')),
]


def build_parser():
    parser = argparse.ArgumentParser()
    for flags, kw in CLI:
        parser.add_argument(*flags, **kw)
    return parser


def main(argv=None):
    signal.signal(signal.SIGTERM, signal_handler)
    signal.signal(signal.SIGINT, signal_handler)
    args = vars(build_parser().parse_args(argv))
    seed = dropin.time_seed() if args['rand_seed'] < 0 else args['rand_seed']   # :415-419
    np.random.seed(seed)
    random.seed(seed)
    try:
        import nxsdk  # noqa: F401
    except ImportError:
        if args['run_mode'] == 'loihi' or args['fugu_backend'] == 'loihi':
            raise SystemExit('\n *** NxSDK package is not installed. Running RandomWalk on Loihi is unavailable. *** \n')
    log_file = os.path.expanduser(args['log_file'])
    os.makedirs(os.path.dirname(log_file), exist_ok=True)        # (several ranks may start at once)
    utils.setup_logging(log_file=log_file, useLogFile=True, log_level=args['loglevel'])
    rw_logger.info('Using config: {}'.format(args))
    rw_logger.info('Starting run on {}'.format(datetime.now()))
    run(args)
    rw_logger.info('Completed run on {}'.format(datetime.now()))


if __name__ == "__main__":
    main()
