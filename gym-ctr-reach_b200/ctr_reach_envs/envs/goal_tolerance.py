"""Goal-tolerance curriculum (host-side scalar schedule; the value is pushed to the device with ctr_set_tolerance).
Semantics of the reference's GoalTolerance (envs/goal_tolerance.py:8-55), including the quirk that the 'constant'
function returns `initial_tol` until the first update() and `final_tol` afterwards."""
import numpy as np


class GoalTolerance(object):
    def __init__(self, goal_tolerance_parameters):
        self.goal_tolerance_parameters = goal_tolerance_parameters
        p = goal_tolerance_parameters
        self.inc_tol_obs = p['inc_tol_obs']
        self.init_tol, self.final_tol, self.N_ts, self.function = p['initial_tol'], p['final_tol'], p['N_ts'], p['function']
        assert self.function in ['constant', 'linear', 'decay'], 'Not a valid function. Choose constant, linear or decay.'
        if self.function == 'linear':
            self.a = (self.final_tol - self.init_tol) / self.N_ts
            self.b = self.init_tol
        if self.function == 'decay':
            self.a = self.init_tol
            self.r = 1 - np.power((self.final_tol / self.init_tol), 1 / self.N_ts)
        self.set_tol_value = p['set_tol']
        self.current_tol = self.init_tol if self.set_tol_value == 0 else self.set_tol_value
        self.training_step = 0

    def update(self, timestep):
        if self.set_tol_value != 0:
            self.current_tol = self.set_tol_value
        elif self.function == 'linear' and timestep <= self.N_ts:
            self.current_tol = self.a * timestep + self.b
        elif self.function == 'decay' and timestep <= self.N_ts:
            self.current_tol = self.a * np.power(1 - self.r, timestep)
        else:
            self.current_tol = self.final_tol

    def get_tol(self):
        return self.current_tol
