"""TEST INFRASTRUCTURE: a deterministic stand-in for ``ortools.graph.pywrapgraph.SimpleMinCostFlow``
(ortools is not installable here) with the methods libmot's MinCostFlowTracker calls
(libmot/libmot/data_association/offline_mincostflow.py:88,157-162,197-198,351-389): AddArcWithCapacityAndUnitCost,
SetNodeSupply, Solve, OPTIMAL, OptimalCost, NumArcs, Tail, Head, Flow.

Successive shortest paths (Bellman-Ford on the residual graph, first-found predecessor wins, arcs scanned in
insertion order), integer costs.  It returns A minimum-cost flow; with several optimal flows it need not pick
the one ortools picks -- so it pins that two arms fed the same arcs in the same order get the same IDs, and
lets the reference's whole _assign_IDs_to_detections -> compute_trajectories -> libmot_det2det chain run; it
does not pin ortools' tie-breaking."""


class SimpleMinCostFlow:
    OPTIMAL = 1
    INFEASIBLE = 3

    def __init__(self):
        self._tail, self._head, self._cap, self._cost = [], [], [], []
        self._supply = {}
        self._flow = []
        self._opt = 0
        self.calls = []              # every AddArc call, for input-identity checks

    # ---- graph construction
    def AddArcWithCapacityAndUnitCost(self, tail, head, capacity, unit_cost):
        self._tail.append(int(tail)); self._head.append(int(head))
        self._cap.append(int(capacity)); self._cost.append(int(unit_cost))
        self.calls.append((int(tail), int(head), int(capacity), int(unit_cost)))
        return len(self._tail) - 1

    def SetNodeSupply(self, node, supply):
        self._supply[int(node)] = int(supply)

    def NumArcs(self):
        return len(self._tail)

    def Tail(self, a):
        return self._tail[a]

    def Head(self, a):
        return self._head[a]

    def Flow(self, a):
        return self._flow[a]

    def OptimalCost(self):
        return self._opt

    # ---- solve: ship `supply` units from the (single) source to the (single) sink
    def Solve(self):
        srcs = [n for n, s in self._supply.items() if s > 0]
        snks = [n for n, s in self._supply.items() if s < 0]
        m = len(self._tail)
        self._flow = [0] * m
        self._opt = 0
        if not srcs and not snks:
            return self.OPTIMAL
        assert len(srcs) == 1 and len(snks) == 1 and self._supply[srcs[0]] == -self._supply[snks[0]]
        s, t, need = srcs[0], snks[0], self._supply[srcs[0]]
        n = 1 + max(max(self._tail, default=0), max(self._head, default=0), s, t)
        INF = float('inf')
        for _ in range(need):
            dist = [INF] * n
            pred = [None] * n            # (arc, forward?)
            dist[s] = 0
            for _round in range(n):
                changed = False
                for a in range(m):
                    u, v, c = self._tail[a], self._head[a], self._cost[a]
                    if self._flow[a] < self._cap[a] and dist[u] + c < dist[v]:
                        dist[v] = dist[u] + c; pred[v] = (a, True); changed = True
                    if self._flow[a] > 0 and dist[v] - c < dist[u]:
                        dist[u] = dist[v] - c; pred[u] = (a, False); changed = True
                if not changed:
                    break
            if dist[t] == INF:
                return self.INFEASIBLE
            v = t
            while v != s:
                a, fwd = pred[v]
                if fwd:
                    self._flow[a] += 1; v = self._tail[a]
                else:
                    self._flow[a] -= 1; v = self._head[a]
            self._opt += dist[t]
        return self.OPTIMAL
