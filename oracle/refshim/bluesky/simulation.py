"""ORACLE (test infrastructure, not product code).

Restatement of the slice of ``bluesky/simulation`` the reference envs touch:
``ScreenIO`` (subclassed by bluesky_gym/envs/common/screen_dummy.py:1-9) and
``Simulation.step`` (stack.process -> traf.update -> simt += simdt; SURVEY.md
A.3 last paragraph).  Simplification, documented: a ``DT`` command takes effect
in the same step it is processed.  parity unpinned vs upstream.
"""


class ScreenIO:
    def echo(self, text='', flags=0):
        pass


class Simulation:
    def __init__(self, settings):
        self.settings = settings
        self.simt = 0.0
        self.simdt = settings.simdt
        self.ffmode = False
        self.traf = None
        self.stack = None

    def reset(self):
        self.simt = 0.0

    def step(self):
        self.stack.process()
        if self.traf.ntraf > 0:
            self.traf.update()
            self.simt += self.simdt
