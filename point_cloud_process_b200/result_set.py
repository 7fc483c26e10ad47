"""Result containers with the interface of Kdtree_Octree/lesson2/result_set.py.

    DistIndex(distance, index)            result_set.py:6-13
    KNNResultSet(capacity)                result_set.py:16-60
    RadiusNNResultSet(radius)             result_set.py:63-93

The GPU search entry points (kdtree.py / octree.py of this package) fill these objects in one
shot through `_load()`; `add_point` keeps the reference's incremental semantics for callers that
feed candidates themselves (host bookkeeping only -- no distance is computed here).
"""
import bisect

EMPTY_DIST = 1e10          # result_set.py:19 -- worst_dist of a fresh KNNResultSet, and the value of unfilled slots


class DistIndex:
    __slots__ = ("distance", "index")

    def __init__(self, distance, index):
        self.distance, self.index = distance, index

    def __lt__(self, other):
        return self.distance < other.distance

    def __repr__(self):
        return f"DistIndex({self.distance!r}, {self.index!r})"


class KNNResultSet:
    """Bounded ascending list of the `capacity` nearest (distance, index) pairs."""

    def __init__(self, capacity):
        self.capacity = capacity
        self.count = 0
        self.worst_dist = EMPTY_DIST
        self.dist_index_list = [DistIndex(EMPTY_DIST, 0) for _ in range(capacity)]
        self.comparison_counter = 0

    def size(self):
        return self.count

    def full(self):
        return self.count == self.capacity

    def worstDist(self):
        return self.worst_dist

    def add_point(self, dist, index):
        """result_set.py:35-53: reject iff dist > worst; a candidate equal to existing distances goes
        BEHIND them, and when the list is full it displaces the last slot (last arrival wins a tie
        at the k-th place)."""
        self.comparison_counter += 1
        if dist > self.worst_dist:
            return
        if self.count < self.capacity:
            self.count += 1
        slots = self.dist_index_list
        live = [d.distance for d in slots[:self.count - 1]]
        slots.insert(bisect.bisect_right(live, dist), DistIndex(dist, index))
        del slots[self.count]                       # the displaced (or still unfilled) slot
        self.worst_dist = slots[self.capacity - 1].distance

    def _load(self, dist, idx, count, comparisons=0):
        """Take one row of a batched GPU search: `dist`/`idx` of length capacity, ascending
        (distance, index), slots >= count already (1e10, 0) as result_set.py:19-22 leaves them."""
        count = int(count)
        row = [DistIndex(d, int(i)) for d, i in zip(dist[:count], idx[:count])]
        if self.count:
            # the set already holds candidates (earlier add_point calls or an earlier search on the same object):
            # the reference accumulates through add_point (result_set.py:35-53) -- feed the row through it, in
            # its (distance, index) order, so that ties behave as they do there
            counter = self.comparison_counter
            for d in row:
                self.add_point(d.distance, d.index)
            self.comparison_counter = counter + int(comparisons)
            return
        self.count = len(row)
        self.dist_index_list = row + [DistIndex(EMPTY_DIST, 0) for _ in range(self.capacity - len(row))]
        self.worst_dist = self.dist_index_list[self.capacity - 1].distance if self.capacity else EMPTY_DIST
        self.comparison_counter += int(comparisons)

    def __str__(self):
        rows = ["%d - %.2f\n" % (d.index, d.distance) for d in self.dist_index_list]
        return "".join(rows) + "In total %d comparison operations." % self.comparison_counter


class RadiusNNResultSet:
    """Every (distance, index) with NOT distance > radius (result_set.py:78-84)."""

    def __init__(self, radius):
        self.radius = radius
        self.count = 0
        self.worst_dist = radius
        self.dist_index_list = []
        self.comparison_counter = 0

    def size(self):
        return self.count

    def worstDist(self):
        return self.radius

    def add_point(self, dist, index):
        self.comparison_counter += 1
        if dist > self.radius:
            return
        self.count += 1
        self.dist_index_list.append(DistIndex(dist, index))

    def _load(self, dist, idx, comparisons=0):
        """Append one CSR segment of a GPU radius search (sorted by (distance, index))."""
        self.dist_index_list.extend(DistIndex(d, int(i)) for d, i in zip(dist, idx))
        self.count += len(idx)
        self.comparison_counter += int(comparisons)

    def __str__(self):
        self.dist_index_list.sort()
        rows = ["%d - %.2f\n" % (d.index, d.distance) for d in self.dist_index_list]
        return "".join(rows) + ("In total %d neighbors within %f.\nThere are %d comparison operations."
                                % (self.count, self.radius, self.comparison_counter))
