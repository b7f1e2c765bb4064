"""ORACLE (test infrastructure, not product code).

Restatement of the few stack commands the reference envs queue
(horizontal_cr_env.py:72,275; sector_cr_env.py:77,321-322; merge_env.py:157):
DT, FF, HDG, SPD, plus ASAS ON / RESO MVP / RESOOFF used by the north-star
configs.  Commands are queued by ``stack()`` and executed by ``process()`` at the
start of the next ``sim.step()``.  Speed arguments follow BlueSky's ``spd``
parser: values > 1 are knots CAS, converted with ``kts``.  parity unpinned.
"""
from .tools.aero import kts


class Stack:
    def __init__(self, sim, traf):
        self.sim = sim
        self.traf = traf
        self.queue = []

    def stack(self, cmdline):
        for c in cmdline.split(';'):
            c = c.strip()
            if c:
                self.queue.append(c)

    def process(self):
        q, self.queue = self.queue, []
        for line in q:
            self._exec(line)

    def _exec(self, line):
        tok = line.replace(',', ' ').split()
        cmd = tok[0].upper()
        t = self.traf
        # "ACID CMD args" form
        if cmd not in ('DT', 'FF', 'HDG', 'SPD', 'ASAS', 'RESO', 'RESOOFF', 'NORESO', 'RMETHH', 'RMETHV') and len(tok) > 1:
            tok[0], tok[1] = tok[1], tok[0]
            cmd = tok[0].upper()
        if cmd == 'DT':
            self.sim.simdt = float(tok[1])
        elif cmd == 'FF':
            self.sim.ffmode = True
        elif cmd == 'HDG':
            idx = t.id2idx(tok[1])
            if idx >= 0:
                t.ap.selhdgcmd(idx, float(tok[2]))
        elif cmd == 'SPD':
            idx = t.id2idx(tok[1])
            if idx >= 0:
                v = float(tok[2])
                t.ap.selspdcmd(idx, v * kts if abs(v) > 1.0 else v)
        elif cmd == 'ASAS':
            t.cd.enabled = tok[1].upper() == 'ON'
            if not t.cd.enabled:
                t.cd.clear()
        elif cmd == 'RESO':
            t.cr.enabled = tok[1].upper() == 'MVP'        # 'reso off' (merge_env.py:157) switches it off
            if not t.cr.enabled:
                t.cr.active[:] = False
                t.cr.resopairs = set()
        elif cmd == 'RMETHH':     # resolution method, horizontal: BOTH / ON (default) | OFF / NONE
            on = tok[1].upper() in ('BOTH', 'ON')
            t.cr.swresohoriz = on
            if on:
                t.cr.swresovert = False
        elif cmd == 'RMETHV':     # resolution method, vertical: V/S / ON = vertical only | OFF / NONE
            on = tok[1].upper() in ('V/S', 'ON')
            t.cr.swresovert = on
            if on:
                t.cr.swresohoriz = False
        elif cmd == 'RESOOFF':
            idx = t.id2idx(tok[1])
            if idx >= 0:
                t.cr.resooffac[idx] = True
        elif cmd == 'NORESO':
            idx = t.id2idx(tok[1])
            if idx >= 0:
                t.cr.noresoac[idx] = True
        elif cmd in ('ADDWPT', 'DEST'):   # "ACID ADDWPT lat lon" / "ACID DEST lat lon" (merge_env.py:155-156)
            idx = t.id2idx(tok[1])
            if idx >= 0:
                t.ap.addwpt(idx, float(tok[2]), float(tok[3]), dest=(cmd == 'DEST'))
        # anything else is outside the hot path: ignored
