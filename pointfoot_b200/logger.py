"""``Logger`` -- legged_gym/utils/logger.py:36-133 for the headless path: the play loop's state /
reward recorder.  Same data-side API (``log_state``, ``log_states``, ``log_rewards``, ``reset``, ``print_rewards``,
``plot_states``); ``plot_states`` renders the reference's 3 x 3 panels to a PNG through matplotlib's Agg canvas when matplotlib
is installed (no window: there is no display on the box) and always writes the recorded series to ``<stem>.npz``."""
from collections import defaultdict

import numpy as np

# (row, col, title, x label, y label, [(key, legend)]) of the reference figure (logger.py:66-127); "xy" panels plot one key against another
PANELS = (
    (0, 0, "Base velocity x", "time [s]", "base lin vel [m/s]", (("base_vel_x", "measured"), ("command_x", "commanded"))),
    (0, 1, "Base velocity y", "time [s]", "base lin vel [m/s]", (("base_vel_y", "measured"), ("command_y", "commanded"))),
    (0, 2, "Base velocity yaw", "time [s]", "base ang vel [rad/s]", (("base_vel_yaw", "measured"), ("command_yaw", "commanded"))),
    (1, 0, "DOF Position", "time [s]", "Position [rad]", (("dof_pos", "measured"), ("dof_pos_target", "target"))),
    (1, 1, "Joint Velocity", "time [s]", "Velocity [rad/s]", (("dof_vel", "measured"), ("dof_vel_target", "target"))),
    (1, 2, "Base velocity z", "time [s]", "base lin vel [m/s]", (("base_vel_z", "measured"),)),
    (2, 0, "Vertical Contact forces", "time [s]", "Forces z [N]", (("contact_forces_z", "force"),)),
    (2, 2, "Torque", "time [s]", "Joint Torque [Nm]", (("dof_torque", "measured"),)),
)


def _scalar(v):
    return v.item() if hasattr(v, "item") else float(v)


class Logger:
    def __init__(self, dt):
        self.state_log = defaultdict(list)
        self.rew_log = defaultdict(list)
        self.dt = dt
        self.num_episodes = 0
        self.plot_process = None          # (the reference forks a matplotlib window here; nothing to fork headless)

    def log_state(self, key, value):
        self.state_log[key].append(value)

    def log_states(self, dict):
        for key, value in dict.items():
            self.log_state(key, value)

    def log_rewards(self, dict, num_episodes):
        for key, value in dict.items():
            if 'rew' in key:
                self.rew_log[key].append(_scalar(value) * num_episodes)
        self.num_episodes += num_episodes

    def reset(self):
        self.state_log.clear()
        self.rew_log.clear()

    def series(self):
        """recorded states as numpy arrays + the time axis of logger.py:60-62"""
        out = {k: np.asarray([np.asarray(_scalar(x) if np.ndim(x) == 0 else x, dtype=np.float64) for x in v]) for k, v in self.state_log.items() if v}
        n = max((len(v) for v in out.values()), default=0)
        out["time"] = np.linspace(0, n * self.dt, n)
        return out

    def plot_states(self, stem="logger_states"):
        s = self.series()
        np.savez(stem + ".npz", **s)
        try:
            import matplotlib
            if getattr(matplotlib, "__file__", None) is None:       # a stand-in module (test harnesses stub it), not the library
                raise ImportError("matplotlib is a stub")
            matplotlib.use("Agg")
            import matplotlib.pyplot as plt
        except ImportError:
            return stem + ".npz"
        fig, axs = plt.subplots(3, 3, figsize=(14, 10))
        for r, c, title, xl, yl, keys in PANELS:
            a = axs[r, c]
            for key, label in keys:
                if key in s:
                    y = s[key]
                    if y.ndim == 2:
                        for i in range(y.shape[1]):
                            a.plot(s["time"][:len(y)], y[:, i], label=f"{label} {i}")
                    else:
                        a.plot(s["time"][:len(y)], y, label=label)
            a.set(xlabel=xl, ylabel=yl, title=title)
            a.legend()
        a = axs[2, 1]                      # torque / velocity curve (logger.py:117-120)
        if "dof_vel" in s and "dof_torque" in s:
            a.plot(s["dof_vel"], s["dof_torque"], 'x', label='measured')
        a.set(xlabel='Joint vel [rad/s]', ylabel='Joint Torque [Nm]', title='Torque/velocity curves')
        a.legend()
        fig.savefig(stem + ".png")
        plt.close(fig)
        return stem + ".png"

    def print_rewards(self):
        print("Average rewards per second:")
        for key, values in self.rew_log.items():
            mean = np.sum(np.array(values)) / self.num_episodes
            print(f" - {key}: {mean}")
        print(f"Total number of episodes: {self.num_episodes}")
