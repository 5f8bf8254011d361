"""``pymp.collision_matrix`` mirror: AllowedCollisionMatrix (``src/collision_matrix.{h,cpp}``).

Host-side only.  The scene compiler evaluates ``get_allowed_collision`` once per object pair and
hoists the verdict into the admitted pair list (a pure function of names, so it equals the
reference's post-narrowphase ``filterCollisions``, ``src/planning_world.cpp:264-274``).
A ``version`` counter lets the world notice edits and recompile the pair list.
"""
from __future__ import annotations

import enum


class AllowedCollision(enum.Enum):
    """``src/collision_matrix.h:12-17``"""

    NEVER = 0
    ALWAYS = 1
    CONDITIONAL = 2


def _is_name(x):
    return isinstance(x, str)


class AllowedCollisionMatrix:
    def __init__(self):
        self._entries: dict[str, dict[str, AllowedCollision]] = {}
        self._defaults: dict[str, AllowedCollision] = {}
        self.version = 0

    # -- entries ------------------------------------------------------------------------------
    def get_entry(self, name1, name2):
        return self._entries.get(name1, {}).get(name2)

    def has_entry(self, name1, name2=None):
        if name2 is None:
            return name1 in self._entries
        return name2 in self._entries.get(name1, {})

    def _set_pair(self, n1, n2, allowed):
        v = AllowedCollision.ALWAYS if allowed else AllowedCollision.NEVER
        self._entries.setdefault(n1, {})[n2] = v
        self._entries.setdefault(n2, {})[n1] = v

    def set_entry(self, *args):
        """Overloads of ``src/collision_matrix.cpp:34-89``: (name1, name2, allowed),
        (name, other_names, allowed), (names1, names2, allowed), (name, allowed), (names, allowed),
        (allowed)."""
        self.version += 1
        if len(args) == 3:
            a, b, allowed = args
            if _is_name(a) and _is_name(b):
                self._set_pair(a, b, allowed)
            elif _is_name(a):
                for o in b:
                    if o != a:
                        self._set_pair(o, a, allowed)
            else:
                for n1 in a:
                    for o in b:
                        if o != n1:
                            self._set_pair(o, n1, allowed)
        elif len(args) == 2:
            a, allowed = args
            names = [a] if _is_name(a) else list(a)
            for name in names:
                for other in list(self._entries.keys()):
                    if other != name:
                        self._set_pair(name, other, allowed)
        elif len(args) == 1:
            v = AllowedCollision.ALWAYS if args[0] else AllowedCollision.NEVER
            for d in self._entries.values():
                for k in d:
                    d[k] = v
        else:
            raise TypeError("set_entry: incompatible arguments")

    def _remove_pair(self, n1, n2):
        for a, b in ((n1, n2), (n2, n1)):
            d = self._entries.get(a)
            if d is not None and b in d:
                del d[b]
                if not d:
                    del self._entries[a]

    def remove_entry(self, *args):
        """``src/collision_matrix.cpp:91-138``"""
        self.version += 1
        if len(args) == 2:
            a, b = args
            if _is_name(a) and _is_name(b):
                self._remove_pair(a, b)
            elif _is_name(a):
                for o in b:
                    if o != a:
                        self._remove_pair(o, a)
            else:
                for n1 in a:
                    for o in b:
                        if o != n1:
                            self._remove_pair(o, n1)
        elif len(args) == 1:
            names = [args[0]] if _is_name(args[0]) else list(args[0])
            for name in names:
                self._entries.pop(name, None)
                for k in list(self._entries.keys()):
                    d = self._entries[k]
                    if name in d:
                        del d[name]
                        if not d:
                            del self._entries[k]
        else:
            raise TypeError("remove_entry: incompatible arguments")

    def __len__(self):
        return len(self._entries)

    # -- defaults -----------------------------------------------------------------------------
    def get_default_entry(self, name1, name2=None):
        """``src/collision_matrix.cpp:140-164``"""
        if name2 is None:
            return self._defaults.get(name1)
        t1, t2 = self._defaults.get(name1), self._defaults.get(name2)
        if t1 is None and t2 is None:
            return None
        if t2 is None:
            return t1
        if t1 is None:
            return t2
        if AllowedCollision.NEVER in (t1, t2):
            return AllowedCollision.NEVER
        if AllowedCollision.CONDITIONAL in (t1, t2):
            return AllowedCollision.CONDITIONAL
        return AllowedCollision.ALWAYS

    def has_default_entry(self, name):
        return name in self._defaults

    def set_default_entry(self, names, allowed):
        self.version += 1
        for n in [names] if _is_name(names) else names:
            self._defaults[n] = AllowedCollision.ALWAYS if allowed else AllowedCollision.NEVER

    def remove_default_entry(self, names):
        self.version += 1
        for n in [names] if _is_name(names) else names:
            self._defaults.pop(n, None)

    def get_allowed_collision(self, name1, name2):
        """``src/collision_matrix.cpp:194-198``: explicit entry wins, else the default-pair rule."""
        t = self.get_entry(name1, name2)
        return t if t is not None else self.get_default_entry(name1, name2)

    def admits(self, name1, name2):
        """True iff a hit between the two names survives ``filterCollisions``
        (``src/planning_world.cpp:269-272``): no verdict, or NEVER."""
        t = self.get_allowed_collision(name1, name2)
        return t is None or t == AllowedCollision.NEVER

    def clear(self):
        self.version += 1
        self._entries.clear()
        self._defaults.clear()

    def get_all_entry_names(self):
        return sorted(set(self._entries.keys()) | set(self._defaults.keys()))

    def __str__(self):
        names = self.get_all_entry_names()
        lines = []
        for i, n in enumerate(names):
            row = []
            for m in names:
                t = self.get_allowed_collision(n, m)
                row.append("-" if t is None else {"NEVER": "0", "ALWAYS": "1", "CONDITIONAL": "?"}[t.name])
            lines.append(f"{i:3d} {n:>24s} | " + " ".join(row))
        return "\n".join(lines)
