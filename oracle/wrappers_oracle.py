"""TEST INFRASTRUCTURE -- NumPy restatement of the reference's wrapper arithmetic (hydrogym/jax/env_core.py).
Only tests/ may import this.  Pinned: tests/test_reference_shim.py drives the reference's own env_core.py wrapper
stack (unmodified, over the NumPy stand-ins of tests/golden/jax_numpy_shim.py) and these functions with the same
scripted environment and compares at 1e-11; the functions below follow the cited lines statement by statement.

Each function maps (previous wrapper state, what the wrapped env returned this step) -> (outputs, new state)."""
import numpy as np


def log_wrapper_step(state, reward, done):
    """E:526-546.  state = dict(episode_returns, episode_lengths, returned_episode_returns,
    returned_episode_lengths, timestep)."""
    new_episode_return = state["episode_returns"] + reward
    new_episode_length = state["episode_lengths"] + 1
    new = dict(
        episode_returns=new_episode_return * (1 - done),
        episode_lengths=new_episode_length * (1 - done),
        returned_episode_returns=state["returned_episode_returns"] * (1 - done) + new_episode_return * done,
        returned_episode_lengths=state["returned_episode_lengths"] * (1 - done) + new_episode_length * done,
        timestep=state["timestep"] + 1,
    )
    info = dict(returned_episode_returns=new["returned_episode_returns"],
                returned_episode_lengths=new["returned_episode_lengths"], timestep=new["timestep"],
                returned_episode=done)
    return info, new


def _running_moments(mean, var, count, batch):
    """E:636-649 / 662-675 / 721-734 (the same nine statements three times in the reference)."""
    batch_mean = np.mean(batch, axis=0)
    batch_var = np.var(batch, axis=0)
    batch_count = batch.shape[0]
    delta = batch_mean - mean
    tot_count = count + batch_count
    new_mean = mean + delta * batch_count / tot_count
    m_a = var * count
    m_b = batch_var * batch_count
    M2 = m_a + m_b + np.square(delta) * count * batch_count / tot_count
    return new_mean, M2 / tot_count, tot_count


def normalize_obs_reset(obs):
    """E:628-657."""
    mean, var, count = _running_moments(np.zeros_like(obs), np.ones_like(obs), 1e-4, obs)
    return (obs - mean) / np.sqrt(var + 1e-8), dict(mean=mean, var=var, count=count)


def normalize_obs_step(state, obs):
    """E:659-692."""
    mean, var, count = _running_moments(state["mean"], state["var"], state["count"], obs)
    return (obs - mean) / np.sqrt(var + 1e-8), dict(mean=mean, var=var, count=count)


def normalize_reward_reset(batch_count):
    """E:709-719."""
    return dict(mean=0.0, var=1.0, count=1e-4, return_val=np.zeros((batch_count,)))


def normalize_reward_step(state, reward, done, gamma):
    """E:721-742 (note: batch_count comes from obs.shape[0] = len(reward))."""
    return_val = state["return_val"] * gamma * (1 - done) + reward
    mean, var, count = _running_moments(state["mean"], state["var"], state["count"], return_val)
    return reward / np.sqrt(var + 1e-8), dict(mean=mean, var=var, count=count, return_val=return_val)
