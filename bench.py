"""Headline benchmark: env-steps/sec of suite tasks through the batched backend.

  python bench.py --gpus N --steps K --warmup W            # this repo, N GPUs
  python bench.py --impl reference --steps K --warmup W    # CPU arm (see below)

One "step" = one `Environment.step` for every environment of the batch: copy of the
step's actions into the images, `n_sub_steps` physics steps in ONE kernel launch
(engine.py:149-166 inside rl/control.py:100-106) and the task's observation/reward
kernel.  Workload at every N: cheetah-run at 16384 envs per GPU (BASELINE.json
configs[2]; weak scaling, no per-step collective).  "extra" carries, measured the same way
in the same run: humanoid-run at 8192 envs per GPU (configs[3]; with its own FP64 roofline,
CPU baseline and ratio), walker-walk (configs[2]), cartpole / acrobot / pendulum swing-up at
4096 envs (configs[1]), quadruped-escape and finger-spin at 8192 envs per GPU (configs[4]:
65536 envs over 8 GPUs), and a CONTACT-RICH variant of cheetah-run and humanoid-run
(actions 0.1 * U(-1, 1): the robots rest on the ground instead of bouncing).

The reference's own engine (MuJoCo 2.00) is a closed binary that is absent here
(SURVEY.md §0.2), so the reference arm and `cpu_baseline` time the CPU oracle port
(oracle/mjstep_oracle.c) on all host threads: kind "port".
"""

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from dm_control_b200.model import FlatModel   # noqa: E402
from dm_control_b200.suite import tasks as T   # noqa: E402

MODELS = os.path.join(ROOT, "dm_control_b200", "suite", "models")
# algorithmic bytes per env-step (state + obs I/O) and the contract
# estimate v0 of fp64 flops per physics sub-step.
ALGO = {"cheetah": dict(bytes=648, flops=2.0e4), "humanoid": dict(bytes=2048, flops=2.3e5),
        "walker": dict(bytes=704, flops=1.8e4), "cartpole": dict(bytes=176, flops=1.0e4),
        "acrobot": dict(bytes=184, flops=1.0e4), "pendulum": dict(bytes=112, flops=1.5e3),
        "finger": dict(bytes=264, flops=1.0e4),
        # quadruped-escape: 8 * (nu + 2 (nq + 2 nv + na + 1) + nobs + 2) with nobs = 101;
        # per-env terrain (161 604 B float32) is resident and touched sparsely: not counted
        "quadruped": dict(bytes=2200, flops=7.0e5)}
# SURVEY.md 8d binding rule: once the oracle's op counter exists its per-config mean
# replaces the estimates (tools/count_ops.py -> profiles/r1_oracle_opcount.json)
FLOPS_SOURCE = {}
FLOPS_BY_TASK = {}
for _file in ("r1_oracle_opcount.json", "r2_oracle_opcount.json"):
  try:
    with open(os.path.join(ROOT, "profiles", _file)) as _f:
      for _name, _rec in json.load(_f).items():
        _dom = _name.split("-", 1)[0]
        FLOPS_BY_TASK[_name] = (_rec["flops_per_substep"],
                                "oracle op counter, %s (%s)" % (_name, _rec["sample"]))
        if _dom in ALGO and _dom not in FLOPS_SOURCE and "@" not in _name:
          ALGO[_dom]["flops"] = _rec["flops_per_substep"]
          FLOPS_SOURCE[_dom] = "oracle op counter, %s (%s)" % (_name, _rec["sample"])
  except (OSError, ValueError, KeyError):
    pass
# `lanes`: execution shape picked from the lane sweep (profiles/r1b_sweep_lanes.jsonl)
_JSON_OUT = None     # the process's original stdout when fd 1 has been pointed at stderr
WORKLOADS = {"cheetah-run": dict(nenv=16384, njmax=64, nconmax=16, settle=200, lanes=8),
             "humanoid-run": dict(nenv=8192, njmax=120, nconmax=40, settle=100, lanes=16),
             "walker-walk": dict(nenv=16384, njmax=48, nconmax=12, settle=50, lanes=8),
             "cartpole-swingup": dict(nenv=4096, njmax=4, nconmax=1, settle=0, lanes=1),
             "acrobot-swingup": dict(nenv=4096, njmax=1, nconmax=1, settle=0, lanes=1),
             "pendulum-swingup": dict(nenv=4096, njmax=1, nconmax=1, settle=0, lanes=1),
             "finger-spin": dict(nenv=8192, njmax=24, nconmax=6, settle=50, lanes=4),
             # BASELINE.json configs[4]: 65536 envs over 8 GPUs = 8192 per GPU
             "quadruped-escape": dict(nenv=8192, njmax=184, nconmax=42, settle=100, lanes=16),
             "quadruped-walk": dict(nenv=8192, njmax=120, nconmax=40, settle=100, lanes=16)}
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the step kernel from the
# committed `ncu --set full` captures (profiles/), keyed by (task, lanes, envs per GPU)
NCU_TRAFFIC = {("cheetah-run", 8, 16384): (375.1e6, "profiles/r2z_static_cheetah_g8.summary.txt"),
               ("cheetah-run", 4, 16384): (345.5e6, "profiles/r1g_cheetah_g4.summary.txt"),
               ("humanoid-run", 16, 8192): (4.317e9, "profiles/r2z_static_humanoid_g16.summary.txt"),
               ("humanoid-run", 8, 8192): (5.531e9, "profiles/r1g_humanoid_g8.summary.txt")}


def load_task(name):
  domain, task = name.split("-", 1)
  tdef = T.REGISTRY[(domain, task)]
  m = FlatModel.load(os.path.join(MODELS, tdef.model + ".npz"))
  return tdef, m


def initial_qpos(m, nenv, seed):
  """Uniform in joint ranges for limited joints (cheetah.py:63-71 style)."""
  rng = np.random.RandomState(seed)
  qpos = np.tile(m.qpos0, (nenv, 1))
  for j in range(m.njnt):
    if m.jnt_type[j] >= 2 and m.jnt_limited[j]:
      lo, hi = m.jnt_range[j]
      qpos[:, m.jnt_qposadr[j]] = rng.uniform(lo, hi, nenv)
  return qpos


def initial_state(name, m, nenv, seed):
  """qpos of the bench batch.  quadruped: the model's standing pose with hinge angles
  spread over the inner half of their ranges, dropped from 0.3 m above its rest height
  with a random heading (random angles at rest height would start 0.3 m inside the
  ground; the task itself starts from a non-contacting height, quadruped.py:245-277)."""
  if not name.startswith("quadruped"):
    return initial_qpos(m, nenv, seed)
  rng = np.random.RandomState(seed)
  qpos = np.tile(m.qpos0, (nenv, 1))
  for j in range(m.njnt):
    if m.jnt_type[j] >= 2 and m.jnt_limited[j]:
      lo, hi = m.jnt_range[j]
      mid, half = 0.5 * (lo + hi), 0.25 * (hi - lo)
      qpos[:, m.jnt_qposadr[j]] = rng.uniform(mid - half, mid + half, nenv)
  yaw = rng.uniform(0, 2 * np.pi, nenv)
  qpos[:, 2] += 0.3
  qpos[:, 3] = np.cos(yaw / 2); qpos[:, 4:6] = 0; qpos[:, 6] = np.sin(yaw / 2)
  return qpos


def initial_hfield(name, m, nenv, seed):
  """quadruped-escape: one terrain per env, drawn as Escape.initialize_episode does
  (quadruped.py:366-382) from RandomState(seed + env); float32 like mjModel.hfield_data."""
  if int(m.nhfielddata) == 0:
    return None
  from dm_control_b200 import suite
  out = np.empty((nenv, int(m.nhfielddata)), dtype=np.float32)
  for e in range(nenv):
    out[e] = suite._escape_terrain(m, np.random.RandomState(seed + e))
  return out


class ClockSampler:
  """nvidia-smi clocks + throttle reasons during the timed region."""
  Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
       "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
       "clocks_event_reasons.sw_power_cap")

  def __init__(self, index):
    self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    try:
      self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                 "--format=csv,noheader,nounits", "-lms", "100"],
                                stdout=self.f, stderr=subprocess.DEVNULL)
    except OSError:
      self.p = None

  def stop(self):
    if self.p is None:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    time.sleep(0.15)
    self.p.terminate(); self.p.wait()
    self.f.flush(); self.f.seek(0)
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in self.f:
      parts = [x.strip() for x in line.split(",")]
      if len(parts) < 6:
        continue
      try:
        sm.append(float(parts[0])); mx.append(float(parts[1]))
      except ValueError:
        continue
      for n, v in zip(names, parts[2:6]):
        if v.lower().startswith("active"):
          reasons.add(n)
    self.f.close(); os.unlink(self.f.name)
    return {"sm_mhz": float(np.median(sm)) if sm else None,
            "sm_max_mhz": float(np.max(mx)) if mx else None, "reasons": sorted(reasons)}


LANES = {"value": 0}


def workload_name(name, nenv, nsub, scale=1.0):
  act = "U(-1,1)" if scale == 1.0 else "%g * U(-1,1)" % scale
  return ("%s, %d envs per GPU, random actions %s, %d physics sub-step(s) per env-step"
          % (name, nenv, act, nsub))


def make_batch(name, nenv, device, seed):
  from dm_control_b200 import binding
  w = WORKLOADS[name]
  tdef, m = load_task(name)
  model = binding.Model(m, device)
  batch = binding.Batch(model, nenv, w["njmax"], w["nconmax"], LANES["value"] or w["lanes"])
  ip, fp = tdef.params(m)
  batch.set_task(tdef.kind, ip, fp)
  batch.write("qpos", initial_state(name, m, nenv, seed))
  hf = initial_hfield(name, m, nenv, seed)
  if hf is not None:
    batch.write_hfield(hf)
  batch.forward(disable_actuation=True)
  nsub = T.n_sub_steps(tdef, m)
  return tdef, m, model, batch, nsub


def run_gpu(name, nenv, steps, warmup, rank, world, device, action_scale=1.0):
  import torch
  from dm_control_b200 import binding
  tdef, m, model, batch, nsub = make_batch(name, nenv, device, seed=1000 + rank)
  stream = torch.cuda.ExternalStream(batch.stream, device=device)
  nu, nobs = m.nu, batch.nobs
  gen = torch.Generator(device="cuda:%d" % device); gen.manual_seed(rank)
  total = steps + warmup
  ctrl = (torch.rand((total, nenv, nu), generator=gen, device="cuda:%d" % device,
                     dtype=torch.float64) * 2 - 1) * action_scale
  obs = torch.empty((nenv, nobs), device="cuda:%d" % device, dtype=torch.float64)
  rew = torch.empty((nenv,), device="cuda:%d" % device, dtype=torch.float64)
  torch.cuda.synchronize(device)

  def env_step(k):          # ONE launch: ctrl scatter + nsub physics steps + obs / reward
    batch.env_step(nsub, ctrl[k].data_ptr(), obs.data_ptr(), rew.data_ptr())

  # settle: the task's own warm-up of the physics (e.g. cheetah.py:73-75), untimed
  settle = WORKLOADS[name]["settle"]
  for k in range(settle):
    batch.copy_from_device("ctrl", ctrl[k % total].data_ptr())
    batch.step(1 if name.startswith("cheetah") else nsub)
  for k in range(warmup):
    env_step(k)
  batch.sync()

  def barrier():
    if world > 1:
      torch.distributed.barrier()
    torch.cuda.synchronize(device)

  # host buffers of the end-to-end section, allocated (pinned) before anything is timed
  rng = np.random.RandomState(7 + rank)
  ew = max(2, warmup)
  h_ctrl = torch.from_numpy(action_scale * rng.uniform(-1, 1, (steps + ew, nenv, nu))).pin_memory()
  h_obs = torch.empty((nenv, nobs), dtype=torch.float64).pin_memory()
  h_rew = torch.empty((nenv,), dtype=torch.float64).pin_memory()
  hc, ho, hr = h_ctrl.numpy(), h_obs.numpy(), h_rew.numpy()

  # ---- device-resident timing (value) ---------------------------------------------
  ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
  kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
         for _ in range(steps)]
  sampler = ClockSampler(device) if rank == 0 else None
  barrier()
  ev[0].record(stream)
  for k in range(steps):
    kev[k][0].record(stream)
    batch.env_step(nsub, ctrl[warmup + k].data_ptr(), obs.data_ptr(), rew.data_ptr())
    kev[k][1].record(stream)
  ev[1].record(stream)
  batch.sync()
  barrier()
  ms_total = ev[0].elapsed_time(ev[1])
  ms_kernel = float(np.mean([a.elapsed_time(b) for a, b in kev]))
  clocks = sampler.stop() if sampler else None

  # ---- end to end through the C ABI with pinned host buffers (e2e) ----------------------
  # every step's actions wait in pinned host memory (one [nenv][nu] buffer per step); the
  # timed region is the public host call: H2D of that buffer, the fused step, D2H of
  # observations and rewards into pinned host buffers, stream synchronisation
  for k in range(ew):
    batch.env_step_host(nsub, hc[k], ho, hr)
  barrier()
  t0 = time.perf_counter()
  _ts = []
  for k in range(steps):
    _t = time.perf_counter()
    batch.env_step_host(nsub, hc[ew + k], ho, hr)
    _ts.append((time.perf_counter() - _t) * 1e6)
  torch.cuda.synchronize(device)
  e2e_s = time.perf_counter() - t0
  if os.environ.get("BENCH_E2E_TRACE"):
    print("e2e per-step us:", " ".join("%.0f" % t for t in _ts), "total %.0f" % (e2e_s * 1e6), file=sys.stderr)
    print("device-resident kernel us:", " ".join("%.0f" % (a.elapsed_time(b) * 1e3) for a, b in kev), file=sys.stderr)
  barrier()
  warn = batch.warnings().sum(axis=0)
  finite = bool(np.all(np.isfinite(ho)) and np.all(np.isfinite(hr)))
  stats = dict(ncon=float(batch.read("ncon").mean()), nefc=float(batch.read("nefc").mean()),
               solver_iter=float(batch.read("solver_iter").mean()),
               reward_mean=float(hr.mean()))
  out = dict(ms_total=ms_total, ms_kernel=ms_kernel, e2e_s=e2e_s, nsub=nsub, nu=nu,
             nobs=nobs, clocks=clocks, warnings=[int(x) for x in warn], finite=finite,
             stats=stats, smem=batch.smem_bytes, stride=batch.stride, lanes=batch.lanes)
  batch.free(); model.free()
  return out


def cpu_run(name, nenv, steps, threads, seed=0, action_scale=1.0):
  """Times the oracle port on host threads: `steps` env-steps of `nenv` envs."""
  from oracle import oracle_py
  tdef, m = load_task(name)
  nsub = T.n_sub_steps(tdef, m)
  om = oracle_py.OracleModel(m)
  rng = np.random.RandomState(seed)
  qpos = initial_state(name, m, nenv, seed)
  qvel = np.zeros((nenv, m.nv)); act = np.zeros((nenv, max(m.na, 1)))
  warm = np.zeros((nenv, m.nv)); tm = np.zeros(nenv)
  ctrl = action_scale * rng.uniform(-1, 1, (steps, nenv, m.nu))
  hf = initial_hfield(name, m, nenv, seed)
  t0 = time.perf_counter()
  oracle_py.run_batch(om, qpos, qvel, act, warm, tm, ctrl, steps, nsub, threads, hfield=hf)
  dt = time.perf_counter() - t0
  return nenv * steps / dt, dt, nsub


def cpu_baseline(name, budget_s=15.0):
  """~budget_s seconds of oracle work on every host thread: chunks of <= 1000 env-steps
  of a 16-envs-per-thread batch (each chunk restarts from a fresh random state)."""
  threads = os.cpu_count() or 1
  nenv = 16 * threads
  v, dt, nsub = cpu_run(name, nenv, 20, threads)               # calibration
  chunk = int(max(20, min(1000, 1.0 * v / nenv)))              # ~1 s per chunk
  done, spent, k = 0, 0.0, 0
  while spent < budget_s:
    _, dt, _ = cpu_run(name, nenv, chunk, threads, seed=1 + k)
    done += nenv * chunk; spent += dt; k += 1
  return {"value": done / spent, "unit": "env-steps/s", "cores": threads, "kind": "port",
          "sample": "%s: %d envs x %d env-steps in %d chunks (x%d physics sub-steps each), "
                    "%d pthreads, %.1f s of oracle time" % (name, nenv, chunk * k, k, nsub,
                                                            threads, spent)}


def _vecenv_worker(conn, name, seed):
  """One process = one environment, as scripts/vec_env.py:346-394 (`_subproc_worker`): waits
  for an action, takes one Environment.step (n_sub_steps physics steps, observation, reward
  -- the Python of rl/control.py:94-123 around the engine calls), sends the TimeStep back."""
  from oracle import oracle_py, task_oracle
  tdef, m = load_task(name)
  nsub = T.n_sub_steps(tdef, m)
  o = oracle_py.Oracle(oracle_py.OracleModel(m))
  o.reset()
  o.qpos[:] = initial_state(name, m, 1, seed)[0]
  o.forward()
  fields = ("qpos", "qvel", "ctrl", "xmat", "xpos", "xquat", "site_xpos", "site_xmat",
            "sensordata", "subtree_com", "geom_xpos", "act", "cvel", "xipos")
  _, fp = tdef.params(m)
  speed = fp[0] if tdef.domain in ("walker", "humanoid", "quadruped", "humanoid_CMU") else None
  while True:
    a = conn.recv()
    if a is None:
      break
    o.ctrl[:] = a
    for _ in range(nsub):
      o.step()
    d = {f: np.array(getattr(o, f), dtype=np.float64).reshape(1, -1) for f in fields}
    obs, r = task_oracle.obs_reward(tdef.domain, tdef.task, m, d, speed)
    conn.send((obs[0], float(r[0])))
  conn.close()


def cpu_python_vecenv(name, budget_s=6.0):
  """The number a dm_control user gets on the host (SURVEY.md 8d second CPU figure): the
  Python Environment.step loop in the SubprocVecEnv shape of scripts/vec_env.py:396-459 --
  one process per core, one env per process, actions and TimeSteps through pipes,
  step_async to all then step_wait on all.  Engine = the CPU oracle port, observation /
  reward = oracle/task_oracle.py (numpy, like the reference's task classes)."""
  import multiprocessing as mp
  ctx = mp.get_context("fork")
  nproc = os.cpu_count() or 1
  tdef, m = load_task(name)
  pipes, procs = [], []
  for i in range(nproc):
    a, b = ctx.Pipe()
    pr = ctx.Process(target=_vecenv_worker, args=(b, name, 100 + i), daemon=True)
    pr.start(); b.close()
    pipes.append(a); procs.append(pr)
  rng = np.random.RandomState(0)

  def step():
    acts = rng.uniform(-1, 1, (nproc, m.nu))
    for c, a in zip(pipes, acts):          # step_async
      c.send(a)
    return [c.recv() for c in pipes]       # step_wait

  for _ in range(5):
    step()
  n, t0 = 0, time.perf_counter()
  while time.perf_counter() - t0 < budget_s:
    step(); n += 1
  dt = time.perf_counter() - t0
  for c in pipes:
    c.send(None)
  for pr in procs:
    pr.join(timeout=5)
  return {"value": nproc * n / dt, "unit": "env-steps/s", "cores": nproc, "kind": "port",
          "shape": "SubprocVecEnv: one Python process per core, one env per process, pipes "
                   "(scripts/vec_env.py:396-459)",
          "sample": "%s: %d processes x %d env-steps, %.1f s" % (name, nproc, n, dt)}


def reference_arm(args):
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  name = args.task
  threads = os.cpu_count() or 1
  nenv = 16 * threads
  # size one "step" (= one env-step of the sample batch) so the run stays short
  v, _, nsub = cpu_run(name, nenv, 20, threads)
  per_step = nenv / v
  for _ in range(args.warmup):
    cpu_run(name, nenv, 1, threads)
  reps = max(1, int(0.5 / max(per_step, 1e-6)))               # >= 0.5 s per timed step
  dt = 0.0
  for k in range(args.steps):
    dt += cpu_run(name, nenv, reps, threads, seed=k)[1]         # oracle time only
  value = nenv * reps * args.steps / dt
  sample = ("%s: %d envs x %d env-steps per step (x%d sub-steps), %d pthreads"
            % (name, nenv, reps, nsub, threads))
  line = {"impl": "reference", "metric": "env-steps/sec (%s)" % name, "value": value,
          "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
          "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
          "data": "synthetic",
          "config": {"workload": workload_name(name, WORKLOADS[name]["nenv"], nsub),
                     "sample": sample, "note": "reference engine (MuJoCo 2.00) is a "
                     "closed binary absent from the reference tree; CPU oracle port timed"},
          "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads,
                           "kind": "port", "sample": sample},
          "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0,
                  "d2h_bytes_per_step": 0}}
  print(json.dumps(line), file=_JSON_OUT or sys.stdout)
  (_JSON_OUT or sys.stdout).flush()


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=20)
  ap.add_argument("--warmup", type=int, default=5)
  ap.add_argument("--impl", default="b200")
  ap.add_argument("--task", default="cheetah-run")
  ap.add_argument("--nenv", type=int, default=0, help="envs per GPU (0 = workload default)")
  ap.add_argument("--lanes", type=int, default=0, help="GPU lanes per environment")
  ap.add_argument("--no-extra", action="store_true")
  ap.add_argument("--no-cpu", action="store_true")
  args = ap.parse_args()
  args.warmup = max(args.warmup, 3)
  LANES["value"] = args.lanes
  if args.impl == "reference":
    return reference_arm(args)

  import torch
  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  torch.cuda.set_device(local)
  if world > 1:
    # libraries (NCCL's version banner) write to fd 1; the bench's stdout is the one JSON
    # line: keep the real stdout aside and point fd 1 at stderr for everything else
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))

  def reduce_max(x):
    if world == 1:
      return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda:%d" % local)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())

  name = args.task
  nenv = args.nenv or WORKLOADS[name]["nenv"]
  from dm_control_b200 import binding
  peaks = {}
  try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
  except (OSError, ValueError):
    pass
  hbm_peak = peaks.get("hbm_gbs", 6650.0)
  fp64_peak = binding.fp64_peak_tflops(local)

  def measure(task, n_env, scale=1.0):
    """One workload on every rank; returns the whole-job record (max over ranks)."""
    r = run_gpu(task, n_env, args.steps, args.warmup, rank, world, local, scale)
    ms_total = reduce_max(r["ms_total"]); e2e_s = reduce_max(r["e2e_s"])
    ms_kernel = reduce_max(r["ms_kernel"])
    domain = task.split("-")[0]
    key = task if scale == 1.0 else task + "@rest"
    flops_sub, flops_src = FLOPS_BY_TASK.get(
        key, FLOPS_BY_TASK.get(task, (ALGO[domain]["flops"],
                                      FLOPS_SOURCE.get(domain, "SURVEY 8d contract estimate v0"))))
    algo_bytes = (2016 if task == "quadruped-walk" else ALGO[domain]["bytes"]) * n_env
    fl = flops_sub * r["nsub"] * n_env / (ms_kernel * 1e-3) / 1e12
    hb = algo_bytes / (ms_kernel * 1e-3) / 1e9
    traffic, traffic_src = NCU_TRAFFIC.get((task, r["lanes"], n_env), (None, None))
    if scale != 1.0:
      traffic, traffic_src = None, None
    rec = {
        "value": n_env * world * args.steps / (ms_total * 1e-3),
        "e2e": n_env * world * args.steps / e2e_s, "unit": "env-steps/s",
        "workload": workload_name(task, n_env, r["nsub"], scale),
        "envs_per_gpu": n_env, "n_sub_steps": r["nsub"], "lanes_per_env": r["lanes"],
        "ms_per_step": ms_total / args.steps, "step_kernel_ms": ms_kernel,
        "h2d_bytes_per_step": n_env * r["nu"] * 8, "d2h_bytes_per_step": n_env * (r["nobs"] + 1) * 8,
        "warnings": r["warnings"], "finite": r["finite"], "stats": r["stats"],
        "smem_bytes_per_env": r["smem"], "stride": r["stride"], "clocks": r["clocks"],
        # the binding roofline: the FP64 vector pipe
        "roofline": {"bound": "fp64", "achieved": fl, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": fl / fp64_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "bp_step_kernel", "kernel_ms": ms_kernel,
                     "flops_per_substep": flops_sub, "flops_source": flops_src,
                     "peak_source": "DFMA microbenchmark in this run (bp_fp64_peak)"},
        # secondary: algorithmic bytes (state + obs I/O) against the measured HBM copy peak
        "roofline_hbm": {"bound": "hbm", "achieved": hb, "peak": hbm_peak, "unit": "GB/s",
                         "frac": hb / hbm_peak, "algorithmic_bytes_per_launch": algo_bytes,
                         "image_bytes_per_launch": 2 * n_env * r["stride"] * 8,
                         "peak_source": "measured" if peaks else "fallback"}}
    return rec

  main_rec = measure(name, nenv)
  extra = {}
  if not args.no_extra and name == "cheetah-run":
    for task in ("humanoid-run", "walker-walk", "cartpole-swingup", "acrobot-swingup",
                 "pendulum-swingup", "quadruped-escape", "finger-spin"):
      extra[task] = measure(task, WORKLOADS[task]["nenv"])
    # contact-rich variants: small actions, the robots rest on the ground
    for task in ("cheetah-run", "humanoid-run"):
      extra[task + " (contact-rich)"] = measure(task, WORKLOADS[task]["nenv"], 0.1)
  # the design's one collective: all-gather of episode statistics over NCCL (sharding.py)
  gathered = None
  if world > 1:
    from dm_control_b200 import sharding
    st = torch.tensor([[main_rec["stats"]["reward_mean"], float(rank)]], dtype=torch.float64,
                      device="cuda:%d" % local)          # one row of statistics per rank
    g = sharding.gather_episode_stats(st)
    gathered = {"ranks_seen": int(g.shape[0]), "rank_ids": [int(x) for x in g[:, 1].tolist()]}
  if rank != 0:
    if world > 1:
      torch.distributed.destroy_process_group()
    return
  r = main_rec
  line = {
      "metric": "env-steps/sec (%s)" % name, "value": r["value"], "unit": "env-steps/s",
      "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
      "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
      "config": {"workload": r["workload"],
                 "envs_per_gpu": nenv, "n_sub_steps": r["n_sub_steps"],
                 "l2": "per-GPU image %.0f MB > 126 MB L2 (inputs larger than L2)"
                       % (nenv * r["stride"] * 8 / 1e6),
                 "smem_bytes_per_env": r["smem_bytes_per_env"],
                 "lanes_per_env": r["lanes_per_env"]},
      "e2e": {"value": r["e2e"], "unit": "env-steps/s",
              "h2d_bytes_per_step": r["h2d_bytes_per_step"],
              "d2h_bytes_per_step": r["d2h_bytes_per_step"]},
      # per env-step: ONE bp_step_kernel launch (ctrl scatter and obs / reward are fused in)
      "gpu_launches": args.steps * world,
      "clocks": r["clocks"],
      "roofline": r["roofline"], "roofline_hbm": r["roofline_hbm"],
      "physics_steps_per_s": r["value"] * r["n_sub_steps"],
      "warnings": r["warnings"], "finite": r["finite"], "stats": r["stats"],
      "extra": extra,
  }
  if gathered is not None:
    line["episode_stats_allgather"] = gathered
  if world == 1 and not args.no_cpu:
    line["cpu_baseline"] = cpu_baseline(name)
    line["cpu_baseline_python_vecenv"] = cpu_python_vecenv(name)
    line["vs_cpu"] = {"e2e_over_c_loop": line["e2e"]["value"] / line["cpu_baseline"]["value"],
                      "e2e_over_python_vecenv":
                          line["e2e"]["value"] / line["cpu_baseline_python_vecenv"]["value"]}
    if "humanoid-run" in extra:      # the other half of the metric gets the same anchors
      h = extra["humanoid-run"]
      h["cpu_baseline"] = cpu_baseline("humanoid-run", budget_s=12.0)
      h["cpu_baseline_python_vecenv"] = cpu_python_vecenv("humanoid-run", budget_s=5.0)
      h["vs_cpu"] = {"e2e_over_c_loop": h["e2e"] / h["cpu_baseline"]["value"],
                     "e2e_over_python_vecenv":
                         h["e2e"] / h["cpu_baseline_python_vecenv"]["value"]}
  print(json.dumps(line), file=_JSON_OUT or sys.stdout)
  (_JSON_OUT or sys.stdout).flush()
  if world > 1:
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
  main()
