"""Angular-fluence miniapp -- drop-in for the reference's ``fluence_mini_app.py`` on a B200.

Same entry points (``run(args)``, ``run_pc``, ``run_functions``, ``sourceTerm``, ``plot_results``), same
command line (README.md:76-134 of the reference), same output files.  ``--run_mode pc`` runs on the
GPU through ``nrw`` (libnrw.so); the neuromorphic modes (fugu / loihi / spinnaker) are dispatch-table
entries that explain what is missing.  Additive flags: ``--rng``, ``--source``, ``--sampler``.

Reference lines are cited as fluence_mini_app.py:<line>.
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

os.environ['SLURM'] = "1"                 # fluence_mini_app.py:7 (Q16: NxSDK habit, harmless)

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import utils.logging_utils as utils       # noqa: E402
from nrw import dist as ndist, dropin, mesh   # noqa: E402
from nrw.fluence import solve_fluence      # noqa: E402

rw_logger = logging.getLogger("NRW")


def signal_handler(signum, frame):
    """fluence_mini_app.py:20-23."""
    rw_logger.info('Signal handler called with signal {}'.format(signum))
    os.killpg(os.getpid(), signum)
    sys.exit(0)


def sourceTerm(x, y):
    """The static source S(x, mu) of fluence_mini_app.py:30-41 (3.0 inside x^2 + y + (7/16)^2 <= 1/4)."""
    return mesh.source_default(x, y)


# ---- back ends: one signature, one return meaning (fluence_mini_app.py:43-249) -----------------

def run_pc(matrix_filename=None, neuraltime=None, initial_walkers=None, meshSize=None,
           runtime_args={}):
    """Conventional-compute back end (fluence_mini_app.py:218-247) on the GPU.

    Returns float64 ``tempdata[meshSize+1, neuraltime]``: walkers on each node per model step, last
    row = walkers that left through the sink AT that step.  ``runtime_args`` may carry
    ``rng`` ('auto' | 'replay' | 'philox') and ``sampler``; the reference's keys are ignored as the
    reference ignores them (Q14).  In replay mode the result and the position of the global
    ``np.random`` stream afterwards are bit-identical to the reference's."""
    csr = mesh.load_matrix(matrix_filename, n_live=meshSize)
    return dropin.pc_tempdata(csr, neuraltime, initial_walkers, meshSize + 1,
                              rng=runtime_args.get('rng'),
                              sampler=runtime_args.get('sampler', 'alias'))


# The 'fugu' and 'loihi' entries keep the reference's adapters (fluence_mini_app.py:43-192) but sit
# on this build's densityrw / loihirw surfaces, i.e. the walk runs on the GPU and is written out as
# the spike table / spike probes those adapters decode.  (The command line still refuses
# ``-r loihi`` without NxSDK, as the reference does, :527-533.)
run_fugu = dropin.fugu_adapter
run_loihi = dropin.loihi_adapter
run_spinnaker = dropin.unavailable_backend('spinnaker', 'sPyNNaker and a SpiNNaker board')

run_functions = {'fugu': run_fugu, 'loihi': run_loihi, 'spinnaker': run_spinnaker, 'pc': run_pc}


def run(args):
    """fluence_mini_app.py:251-447: build the transition matrix, walk M walkers from every mesh
    node, assemble and save the solution."""
    rmode = args['run_mode']
    M = args['num_walks']
    dt, da, v = args['timestep'], args['anglestep'], args['velocity']
    SigS, SigA, L = args['SigS'], args['SigA'], args['interval_length']
    neuraltime = args['neural_timesteps']
    matrix_filename = args['matrix_filename']
    results_filename = args['results_filename']
    source = mesh.SOURCES[args.get('source', 'default')] if args.get('source', 'default') != 'default' \
        else sourceTerm

    global last_run_split
    split = {}
    t_run = time.perf_counter()
    rank, world = ndist.init_from_env()        # under torchrun: one process per GPU, rank 0 writes the files
    try:
        csr = mesh.build_fluence_csr_cached(dt, da, v, SigS, L)       # :279-381 (memoised by the parameters)
    except ValueError:
        rw_logger.error("The probability of more than one scattering event occurring in a single "
                        "time step is greater than 0.05.")
        rw_logger.error("Alter your choice of v, SigS, or dt so that "
                        "1-v*SigS*dt*exp(-v*SigS*dt)-exp(-v*SigS*dt)<=0.05.")
        raise
    m = csr.meta['mesh']
    meshSize, NS, ND = m['meshSize'], m['NS'], m['ND']
    rw_logger.info('Mesh Size: {}'.format(meshSize))
    split['csr_build_s'] = time.perf_counter() - t_run
    if csr.n_rows > mesh.MAX_DENSE_CSV_ROWS:
        matrix_filename = os.path.splitext(matrix_filename)[0] + '.npz'
        rw_logger.warning("matrix has {} rows: saved as CSR in {} instead of a dense CSV".format(
            csr.n_rows, matrix_filename))
    writer = None
    if rank == 0 or rmode != 'pc':
        # :384 (byte-identical text for meshes a dense CSV can hold).  The fused pc path walks the in-memory CSR,
        # so the file is written in the background while the GPU works; every other back end reads it by name
        # and waits for it here
        writer = mesh.MatrixFileWriter(matrix_filename, csr)
        if not (rmode == 'pc' and run_functions['pc'] is run_pc):
            writer.join()
    rw_logger.info("Transition matrix construction complete")
    rw_logger.info("    matrix saved to {}".format(matrix_filename))

    rw_logger.info("Beginning simulation")
    rw_logger.info("    using {} on {}".format(rmode, datetime.now()))
    if rmode == 'pc' and run_functions['pc'] is run_pc:
        # fused path: every start node in one pass, tempdata never materialised
        t0 = time.time()
        res = solve_fluence(csr, M, neuraltime, SigA, dt, v, source=source, rng=args.get('rng'),
                            sampler=args.get('sampler', 'alias'))
        rwSol = res.rwSol
        el = time.time() - t0
        split['solve_s'] = el
        split['walker_steps'] = int(res.walker_steps)
        split['rng'] = res.rng
        split.update(res.split)
        if res.rng == 'replay' and args.get('rng', 'auto') in (None, 'auto'):
            rw_logger.info("    --rng auto chose the bit-exact replay of numpy's stream (one CTA, sequential over start "
                           "nodes, ~2e6 walker-steps/s); --rng philox walks the same mesh at > 1e11 walker-steps/s")
        rw_logger.info('    {} mesh points, {} walker-steps, rng={} in {:.4f} s ({:.3e} walker-steps/s)'
                       .format(meshSize, res.walker_steps, res.rng, el, res.walker_steps / max(el, 1e-12)))
    else:
        # generic plug-in loop of the reference (:397-438) for user-registered back ends
        rwSol = np.zeros((NS - 1, ND - 1))
        run_function = run_functions[rmode]
        tw = dropin.exp_time_weights(SigA, dt, neuraltime)
        for i in range(0, meshSize):
            t0 = time.time()
            tempdata = run_function(matrix_filename=matrix_filename, neuraltime=neuraltime,
                                    initial_walkers=[[i, M]], meshSize=meshSize,
                                    runtime_args={'fugu_backend': args['fugu_backend'],
                                                  'use_sinks': args['use_sinks'],
                                                  'runclean': args['runclean'],
                                                  'rand_seed': args['rand_seed'],
                                                  'rng': args.get('rng'),
                                                  'sampler': args.get('sampler', 'alias')})
            rw_logger.info('    mesh point: {}'.format(i))
            rw_logger.debug('        elapsed time: {}'.format(time.time() - t0))
            posind, aind = mesh.result_index(i, ND)
            tempval = 0
            ks, js = np.nonzero(tempdata[:meshSize, 1:tempdata.shape[1] - 1])
            for j, k in sorted(zip(js + 1, ks)):                      # (t, k) order of :428-434
                tempval = tempval + tempdata[k, j] * tw[j] * v * source(
                    m['meshpoints'][k, 0], m['meshpoints'][k, 1]) * dt
            rwSol[posind, aind] = tempval / M
    rw_logger.info("Simulation using {} completed on {}".format(rmode, datetime.now()))

    if rank == 0:
        if args['plot_results']:
            plot_results(rwSol, m['dirs'], m['space'], results_filename)
        t0 = time.perf_counter()
        mesh.save_array_csv(results_filename, rwSol)                  # :446 (np.savetxt's bytes)
        split['results_file_s'] = time.perf_counter() - t0
        rw_logger.info("Solution saved to {}".format(results_filename))
    if writer is not None:
        t0 = time.perf_counter()
        writer.join()                                                 # the matrix file is complete on return
        split['matrix_file_wait_s'] = time.perf_counter() - t0
        split['matrix_file_skipped_unchanged'] = writer.skipped
    split['total_s'] = time.perf_counter() - t_run
    last_run_split = split
    rw_logger.debug("run() split: {}".format(split))
    return rwSol


last_run_split = {}       # where the time of the most recent run(args) went (seconds; bench.py e2e_run_args)


def plot_results(data, angles, spaces, rfilename):
    """fluence_mini_app.py:449-461 (matplotlib is optional in this build)."""
    try:
        import matplotlib
        matplotlib.use('Agg')
        import matplotlib.pyplot as plt
    except ImportError:
        rw_logger.warning("matplotlib is not installed: skipping the plot")
        return
    rw_logger.info("Plotting solution")
    fig, ax = plt.subplots()
    im = ax.pcolormesh(angles, spaces, data)
    fig.colorbar(im)
    ax.set_title('Angular Fluence')
    ax.set_ylabel('Position')
    ax.set_xlabel('Angle')
    figname = rfilename.split(sep=".")[0] + "_plot.png"
    fig.savefig(figname)
    plt.close(fig)
    rw_logger.info("Plot saved to {}".format(figname))


# ---- command line (fluence_mini_app.py:493-512; README.md:76-134) ------------------------------
_RUN_MODE_HELP = ('The mode specifies how the random walks are ran. Must be one of fugu, loihi, '
                  'spinnaker, or pc. The default is fugu. Specifying pc gives a python implementation.')
CLI = [
    (('-r', '--run_mode'), dict(default='fugu', choices=['fugu', 'loihi', 'spinnaker', 'pc'], help=_RUN_MODE_HELP)),
    (('-M', '--num_walks'), dict(type=int, default=100, help='The number of walkers to start on each location in the mesh.  The default is 100.')),
    (('-dt', '--timestep'), dict(type=float, default=0.01, help='The time step of the simulation. The default is 0.01.')),
    (('-da', '--anglestep'), dict(type=float, default=1.0 / 15.0, help='The angular discretization step size. The default is 1/15.')),
    (('-v', '--velocity'), dict(type=float, default=200, help='The maximum velocity of the particle. The default is 200.')),
    (('-ss', '--SigS'), dict(type=float, default=0.15, help='The scattering cross-section of the particle. This is the rate at which scattering events occur. The default is 0.15.')),
    (('-sa', '--SigA'), dict(type=float, default=0, help='The absorption cross-section.  This is the rate at which particles are absorbed. The default is zero.')),
    (('-L', '--interval_length'), dict(type=float, default=1, help='The length of the spatial interval [-L,L] the particle is allowed to travel through.  The default is 1.')),
    (('-nt', '--neural_timesteps'), dict(type=int, default=30000, help='The number of neural timesteps the simulation is run for. This should be MUCH greater than the number of model time steps desired.  The default is 30000. For pc run mode, this is the number of model timesteps.')),
    (('-d', '--debug'), dict(default=logging.WARNING, action="store_const", dest="loglevel", const=logging.DEBUG, help="Print logging DEBUG statements.")),
    (('--verbose',), dict(action="store_const", dest="loglevel", const=logging.INFO, help="Print logging INFO statements.")),
    (('--matrix_filename',), dict(default='curmat.csv', help='Filename for temporary storage of transition matrix.')),
    (('--results_filename',), dict(default='rwSol.csv', help='Filename to record the results.')),
    (('--fugu_backend',), dict(default='snn', choices=['snn'], help='Runs a spiking simulator using Fugu.')),
    (('--plot_results',), dict(default=False, action='store_true', help='Plots the results after finishing.')),
    (('--log_file',), dict(default='~/.miniapps/fluence_miniapp.log', help='Location of the log file.')),
    (('--runclean',), dict(default=0, help='Run probe-free replicate in a Loihi run.')),
    (('--rand_seed',), dict(type=int, default=-1, help="Sets the seed for the random number generator. Otherwise uses a time-based 'random' RNG seed.")),
    (('--use_sinks',), dict(default=False, action='store_true', help='Include sinks in a fugu simulation; default is False.')),
    # additive (not in the reference)
    (('--rng',), dict(default='auto', choices=list(dropin.RNG_MODES), help="pc mode RNG: 'replay' consumes numpy's legacy stream exactly like the reference (bit-identical results), 'philox' draws on the GPU; 'auto' replays up to ~4M walkers.")),
    (('--source',), dict(default='default', choices=sorted(mesh.SOURCES), help="Source term: 'default' is the script's static source, 'paper' is S=0.015 for |x|<0.5 (Nature Electronics).")),
    (('--sampler',), dict(default='alias', choices=list(dropin.SAMPLERS), help="philox mode transition sampler: O(1) alias tables or the reference's searchsorted on cumulative rows.")),
]


def build_parser():
    parser = argparse.ArgumentParser()
    for flags, kw in CLI:
        parser.add_argument(*flags, **kw)
    return parser


def main(argv=None):
    signal.signal(signal.SIGTERM, signal_handler)                     # :25-26
    signal.signal(signal.SIGINT, signal_handler)
    args = vars(build_parser().parse_args(argv))
    seed = dropin.time_seed() if args['rand_seed'] < 0 else args['rand_seed']   # :517-521
    np.random.seed(seed)                                              # :523-524: seeded ONCE (Q5)
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
