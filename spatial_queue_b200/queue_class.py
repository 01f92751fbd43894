"""Object mirror of the reference's `queue_class.py` on top of the B200 tick and shortest-path library.

Same class names, constructor arguments and attribute names as model/queue_class.py:23-605 (and its Bolinas
copy projects/bolinas_civic/queue_class.py), so a driver keeps its `Network / Node / Link / Agent` objects,
its road-network CSVs, OD CSVs and `special_nodes` JSON.  What changes is where the per-second work runs:

    reference driver (run_traffic_simulation.py:93-131)          this module
    ------------------------------------------------------      ---------------------------------------------
    for agent: get_path / find_next_link / load_vehicle          network.use_gpu(...)   once, after add_demand
    for link:  link.run_link_model(t, ...)                       network.step(t)        every simulated second
    for node:  node.run_node_model(t, ...)                       network.sync_to_host() when the driver reads
                                                                                        object fields

`Link.run_link_model`, `Node.run_node_model` and `Agent.load_vehicle` do not compute on the host: there is no CPU
tick in this package, they raise and point at `Network.step`.  Routing (`Agent.get_path`) goes through
`sp_interface` (the GPU library behind the `sp` ABI), one call per agent like the reference, or in batches via
`Network.route_agents`.

Geometry: the reference projects with geopandas (EPSG:4326 -> EPSG:26910) and measures angles with shapely;
`geo.py` provides the same projection and the two LineString operations the loader needs.  Turn prohibitions
(queue_class.py:69-120, `shapely.ops.split`) are not supported; every shipped configuration leaves them empty.
"""
import json
import math

import numpy as np
import pandas as pd

from . import geo
from . import sp_interface as interface
from ._lib import B200LibraryError
from .flatten import STATUS_CODES, STATUS_NAMES, flatten_network, route_to_links

abs_path = ""   # prefix of the Bolinas-style relative file names (queue_class.py:18 uses the module directory)

_NO_HOST_TICK = ("{} does not run on the host in spatial_queue_b200: call network.use_gpu() once and "
                 "network.step(t) per simulated second (there is no CPU fallback)")


class Node:
    """model/queue_class.py:193-213."""

    def __init__(self, node_id, x, y, node_type, osmid=None):
        self.node_id = node_id
        self.x = x
        self.y = y
        self.node_type = node_type
        self.osmid = osmid
        self.incoming_links = {}    # {in link id: {out link id: turning angle in degrees}}
        self.outgoing_links = []
        self.status = None
        self.node_move = 0
        self.shelter_counts = 0

    def create_virtual_node(self):
        return Node("vn{}".format(self.node_id), self.x + 1, self.y + 1, "v")

    def create_virtual_link(self):
        shape = geo.LineString([(self.x + 1, self.y + 1), (self.x, self.y)])
        return Link("n{}_vl".format(self.node_id), 100, 0, 0, 100000, "v", "vn{}".format(self.node_id),
                    self.node_id, shape)

    def non_conflict_vehicles(self, *a, **k):
        raise B200LibraryError(_NO_HOST_TICK.format("Node.non_conflict_vehicles"))

    def run_node_model(self, *a, **k):
        raise B200LibraryError(_NO_HOST_TICK.format("Node.run_node_model"))


class Link:
    """model/queue_class.py:385-464."""

    def __init__(self, link_id, lanes, length, fft, capacity, link_type, start_nid, end_nid, geometry):
        self.link_id = link_id
        self.lanes = lanes
        self.length = length
        self.fft = fft
        self.capacity = capacity
        self.normal_capacity = capacity
        self.link_type = link_type
        self.start_nid = start_nid
        self.end_nid = end_nid
        self.geometry = geometry
        self.maxmph = length / fft * 2.23694 if fft != 0 else 1e5
        self.storage_capacity = max(18, length * lanes)
        self.remaining_inflow_capacity = capacity / 3600.0
        self.remaining_outflow_capacity = capacity / 3600.0
        self.remaining_storage_capacity = self.storage_capacity
        self.midpoint = geometry.interpolate(0.5, normalized=True).coords[0]
        # heading just after the start and just before the end of the (projected) shape
        sx, sy = geometry.coords[0]
        near_start = geometry.interpolate(0.01, normalized=True)
        self.in_angle = np.arctan2(near_start.y - sy, near_start.x - sx)
        ex, ey = geometry.coords[-1]
        near_end = geometry.interpolate(0.99, normalized=True)
        self.out_angle = np.arctan2(ey - near_end.y, ex - near_end.x)
        self.queue_vehicles = []
        self.run_vehicles = []
        self.travel_time_list = []
        self.travel_time = fft
        self.total_vehicles_left = 0
        self.congested = 0
        self._net = None          # set by Network.use_gpu: status changes are forwarded to the device
        self._status = "open"
        self.burnt_time = 3600 * 100
        self.closed_time = 3600 * 100

    @property
    def status(self):
        return self._status

    @status.setter
    def status(self, value):
        """open / closed / burning / burning_closed / burnt_over (model/dta_meso_butte.py:59-115).  'closed' and
        'burning_closed' zero the inflow capacity in the link model (queue_class.py:444); on the GPU path the change
        takes effect at the next `network.step`."""
        self._status = value
        if self._net is not None:
            self._net._pending_status[self.link_id] = value in ("closed", "burning_closed")

    def run_link_model(self, *a, **k):
        raise B200LibraryError(_NO_HOST_TICK.format("Link.run_link_model"))

    def queue_weight(self):
        """Routing weight of queue_class.py:454 from the (synced) queue length; None when the rule skips the link."""
        if self.status in ("closed", "burning_closed") or self.link_type == "v":
            return None
        return len(self.queue_vehicles) / (self.capacity + 0.1) * 3600 * 25 / self.maxmph + self.fft

    def update_travel_time_by_queue_length(self, g):
        w = self.queue_weight()
        if w is not None:
            g.update_edge(self.start_nid, self.end_nid, w)

    def close_link_to_newcomers(self, g=None):
        g.update_edge(self.start_nid, self.end_nid, 1e8)

    def open_link_to_newcomers(self, g=None):
        g.update_edge(self.start_nid, self.end_nid, self.fft)


class Agent:
    """model/queue_class.py:467-605."""

    def __init__(self, agent_id, origin_nid, destin_nid, departure_time, vehicle_length, gps_reroute, agent_type):
        self.agent_id = agent_id
        self.origin_nid = origin_nid
        self.destin_nid = destin_nid
        self.furthest_nid = destin_nid
        self.departure_time = departure_time
        self.vehicle_length = vehicle_length
        self.gps_reroute = gps_reroute
        self.agent_type = agent_type
        self.current_link_start_nid = "vn{}".format(origin_nid)
        self.current_link_end_nid = origin_nid
        self.route = {}
        self.route_distance = 5e7
        self.fft_distance = 5e7
        self.status = "unloaded"
        self.current_link_enter_time = None
        self.this_link = None
        self.next_link = None
        self.next_link_end_nid = None

    def _adopt(self, t, sp_dist, hops):
        """Shared tail of get_path: the unreachable rules of queue_class.py:494-503, else take the new route."""
        if sp_dist > 1e8:
            if t == 0:
                self.route = {}
                self.status = "shelter_p"
            return False
        self.route_distance = sp_dist
        self.route = {a: b for a, b in hops}
        return True

    def get_path(self, t, g=None):
        sp = g.dijkstra(self.current_link_end_nid, self.destin_nid)
        try:
            d = sp.distance(self.destin_nid)
            self._adopt(t, d, sp.route(self.destin_nid) if d <= 1e8 else ())
        finally:
            sp.clear()

    def get_partial_path(self, t, g=None, link_id_dict=None, node2link_dict=None):
        """Re-plan only the next <= 3 hops and splice (queue_class.py:514-575)."""
        ahead = self.current_link_end_nid
        old_nodes, old_cost = [], 0
        for _ in range(3):
            nxt = self.route.get(ahead)
            if nxt is None:
                break
            old_nodes.append(ahead)
            old_cost += link_id_dict[node2link_dict[(ahead, nxt)]].travel_time
            ahead = nxt
        if ahead == self.current_link_end_nid:
            return None
        sp = g.dijkstra(self.current_link_end_nid, ahead)
        try:
            d = sp.distance(ahead)
            new_cost = 0
            if d > 1e8:
                if t == 0:
                    self.route = {}
                    self.status = "shelter_p"
            else:
                self.route_distance -= old_cost
                for n in old_nodes:
                    del self.route[n]
                for a, b in sp.route(ahead):
                    self.route[a] = b
                    tt = link_id_dict[node2link_dict[(a, b)]].travel_time
                    self.route_distance += tt
                    new_cost += tt
                    if b in self.route:
                        break
        finally:
            sp.clear()
        return d, old_cost, new_cost

    def load_vehicle(self, *a, **k):
        raise B200LibraryError(_NO_HOST_TICK.format("Agent.load_vehicle"))

    def find_next_link(self, node2link_dict=None):
        self.this_link = node2link_dict[(self.current_link_start_nid, self.current_link_end_nid)]
        self.next_link_end_nid = self.route.get(self.current_link_end_nid)
        if self.next_link_end_nid is None:
            self.next_link = None
        else:
            self.next_link = node2link_dict.get((self.current_link_end_nid, self.next_link_end_nid))
            if self.next_link is None:
                self.next_link_end_nid = None

    def move_agent_to_next_link(self, *a, **k):
        raise B200LibraryError(_NO_HOST_TICK.format("Agent.move_agent_to_next_link"))


class Network:
    """model/queue_class.py:23-191 plus the device switch (`use_gpu`, `step`, `run`, `sync_to_host`)."""

    def __init__(self):
        self.nodes = {}
        self.nodes_osmid_dict = {}
        self.links = {}
        self.node2link_dict = {}
        self.g = None
        self.agents = {}
        self.agents_stopped = {}
        self.special_nodes = None
        self.detour_factor = None      # host route mode: shelter_a2 when a reroute lengthens the trip this many times
        self.closure_reroute_freq = None   # host route mode: reroute agents about to enter a closed link, staggered
        self._detoured = []
        self._closed_idx = set()
        self._sim = None
        self._flat = None
        self.route_mode = "device"
        self._pending_status = {}
        self.reroute_freq = 300
        self.weight_freq = 100

    # ---------------------------------------------------------------------------------------------------------
    # loader
    # ---------------------------------------------------------------------------------------------------------
    def dataframe_to_network(self, edges_f=None, nodes_f=None, cf_files=None, special_nodes_f=None, *,
                             project_location=None, network_file_edges=None, network_file_nodes=None,
                             special_nodes=None, scen_nm=None, write_path=None):
        """Accepts both reference signatures: model/queue_class.py:34 (`edges_f, nodes_f, cf_files,
        special_nodes_f`) and the Bolinas copy (`project_location, network_file_edges, network_file_nodes,
        cf_files, special_nodes, scen_nm, write_path`; file names are prefixed with this module's `abs_path`)."""
        bolinas_style = network_file_edges is not None
        prefix = abs_path if bolinas_style else ""
        edges_path = prefix + (network_file_edges if bolinas_style else edges_f)
        nodes_path = prefix + (network_file_nodes if bolinas_style else nodes_f)

        nodes_df = pd.read_csv(nodes_path)
        xy = [geo.utm10n(lon, lat) for lon, lat in zip(nodes_df.lon, nodes_df.lat)]
        nodes_df["x"] = [p[0] for p in xy]
        nodes_df["y"] = [p[1] for p in xy]
        nodes_df = nodes_df[["nid", "x", "y", "osmid"]]

        links_df = pd.read_csv(edges_path)
        for cf_file in cf_files or []:
            cf = pd.read_csv(prefix + cf_file).rename(columns={"lanes": "new_lanes"})
            links_df = links_df.merge(cf[["nid_s", "nid_e", "new_lanes"]], how="left", on=["nid_s", "nid_e"])
            links_df["lanes"] = np.where(np.isnan(links_df["new_lanes"]), links_df["lanes"], links_df["new_lanes"])
        links_df["maxmph"] = np.where(links_df["lanes"] == 0, 0.00000001, links_df["maxmph"])
        links_df["fft"] = links_df["length"] / links_df["maxmph"] * 2.237
        links_df["capacity"] = 1900 * links_df["lanes"]
        links_df["capacity"] = np.where(links_df["capacity"] == 0, 0.01, links_df["capacity"])
        links_df["geometry"] = [geo.project(geo.loads(w)) for w in links_df["geometry"]]
        links_df = links_df[["eid", "nid_s", "nid_e", "type", "lanes", "capacity", "maxmph", "fft", "length", "geometry"]]
        if bolinas_style and write_path is not None and scen_nm is not None:
            try:
                links_df.to_csv(write_path + (project_location or "") +
                                "/simulation_outputs/network/modified_network_edges_{}.csv".format(scen_nm), index=False)
            except OSError:
                pass    # the reference writes this for plotting only

        if special_nodes is None and special_nodes_f is not None:
            with open(special_nodes_f) as fh:
                special_nodes = json.load(fh)
        special_nodes = special_nodes or {"signal": {}, "priority": {}, "prohibition": {}, "closure": []}
        self.special_nodes = special_nodes
        for closed_eid in special_nodes.get("closure", []):
            links_df.loc[links_df["eid"] == closed_eid, "fft"] = 1e8
            links_df.loc[links_df["eid"] == closed_eid, "maxmph"] = 0.00000001
        if len(special_nodes.get("prohibition", {}) or {}) > 0:
            nodes_df, links_df = self._split_prohibition_nodes(nodes_df, links_df, special_nodes["prohibition"])

        self.g = interface.from_dataframe(links_df, "nid_s", "nid_e", "fft")

        # virtual source node + link per real node first, then the real links (queue_class.py:126-145)
        for row in nodes_df.itertuples():
            node = Node(row.nid, row.x, row.y, "real", row.osmid)
            vnode, vlink = node.create_virtual_node(), node.create_virtual_link()
            self.nodes[node.node_id] = node
            self.nodes_osmid_dict[node.osmid] = node.node_id
            self.nodes[vnode.node_id] = vnode
            self.links[vlink.link_id] = vlink
            self.node2link_dict[(vlink.start_nid, vlink.end_nid)] = vlink.link_id
        for row in links_df.itertuples():
            link = Link(row.eid, row.lanes, row.length, row.fft, row.capacity, row.type, row.nid_s, row.nid_e,
                        row.geometry)
            self.links[link.link_id] = link
            key = (link.start_nid, link.end_nid)
            known = self.node2link_dict.get(key)
            if known is None or link.fft < self.links[known].fft:    # parallel links: the faster one is routed on
                self.node2link_dict[key] = link.link_id

    @staticmethod
    def _split_prohibition_nodes(nodes_df, links_df, prohibition):
        """Turn prohibitions (model/queue_class.py:69-120): a node with forbidden turns is taken apart.  Every link
        leaving it starts at a node of its own, placed at 10 % of the link (its geometry keeps what lies beyond a 1 m
        disc around that point), every link entering it ends at a node of its own at 90 %; each allowed
        (in-link, out-link) pair gets a zero-length connector link (100 lanes, capacity 100,000, fft 0) between the
        two.  The old node stays behind without links.  Ids continue the frames' row counts, as in the reference."""
        links_df = links_df.reset_index(drop=True).copy()
        new_nodes, new_links = [], []
        next_node, next_link = nodes_df.shape[0], links_df.shape[0]
        for node_key, turns in prohibition.items():
            node = int(node_key)
            starts, ends = [], []           # (new node id, x, y, link eid) of the out-links / in-links
            for i in links_df.index[links_df["nid_s"] == node]:
                geom, eid = links_df.at[i, "geometry"], links_df.at[i, "eid"]
                pt = geom.interpolate(0.1, normalized=True)
                links_df.at[i, "nid_s"] = next_node
                links_df.at[i, "geometry"] = geo.split(geom, pt.buffer(1))[2]
                starts.append((next_node, pt.x, pt.y, eid))
                next_node += 1
            for i in links_df.index[links_df["nid_e"] == node]:
                geom, eid = links_df.at[i, "geometry"], links_df.at[i, "eid"]
                pt = geom.interpolate(0.9, normalized=True)
                links_df.at[i, "nid_e"] = next_node
                links_df.at[i, "geometry"] = geo.split(geom, pt.buffer(1))[0]
                ends.append((next_node, pt.x, pt.y, eid))
                next_node += 1
            new_nodes += [(n, x, y, "n{}_l{}".format(node, eid)) for n, x, y, eid in starts + ends]
            banned = {(a, b) for a, b in turns}
            for n_in, x_in, y_in, e_in in ends:
                for n_out, x_out, y_out, e_out in starts:
                    if (e_in, e_out) not in banned:
                        new_links.append({"eid": next_link, "nid_s": n_in, "nid_e": n_out, "type": np.nan, "lanes": 100,
                                          "capacity": 100000, "maxmph": 100000, "fft": 0.0, "length": 0.0,
                                          "geometry": geo.LineString([(x_in, y_in), (x_out, y_out)])})
                        next_link += 1
        if new_links:
            links_df = pd.concat([links_df, pd.DataFrame(new_links)[list(links_df.columns)]], ignore_index=True)
        if new_nodes:
            nodes_df = pd.concat([nodes_df, pd.DataFrame(new_nodes, columns=["nid", "x", "y", "osmid"])], ignore_index=True)
        return nodes_df, links_df

    def add_connectivity(self):
        """Per-node turn table (queue_class.py:147-165), wrap rule as written in the reference."""
        for link_id, link in self.links.items():
            self.nodes[link.start_nid].outgoing_links.append(link_id)
            self.nodes[link.end_nid].incoming_links[link_id] = {}
        for node in self.nodes.values():
            for il, table in node.incoming_links.items():
                virtual_in = isinstance(il, str) and il.endswith("vl")
                for ol in node.outgoing_links:
                    turn = 0 if virtual_in else -self.links[il].out_angle + self.links[ol].in_angle
                    if turn >= np.pi:
                        turn = 2 * np.pi - turn
                    if turn <= -np.pi:
                        turn = 2 * np.pi + turn
                    table[ol] = turn / np.pi * 180

    def add_demand(self, ods_fs=None, reroute_pct=0, tow_pct=0, *, demand_files=None, write_path=None):
        """queue_class.py:167-191 (`ods_fs`) and the Bolinas copy (`demand_files`, `write_path`).  Uses numpy's
        global RNG in the reference's order (vehicle length, reroute flag, row shuffle) so that a seeded run
        produces the same agent order."""
        if demand_files is not None:
            prefix = write_path if write_path else abs_path
            files = [prefix + f for f in demand_files]
        else:
            files = list(ods_fs)
        frames = []
        for f in files:
            od = pd.read_csv(f)
            od["origin_nid"] = od["origin_osmid"].map(self.nodes_osmid_dict)
            od["destin_nid"] = od["destin_osmid"].map(self.nodes_osmid_dict)
            if demand_files is None:
                od["dept_time"] = od["dept_time"].astype(int)
            if "agent_id" not in od.columns:
                od["agent_id"] = np.arange(od.shape[0])
            od["veh_len"] = np.random.choice([8, 18], size=od.shape[0], p=[1 - tow_pct, tow_pct])
            od["gps_reroute"] = np.random.choice([0, 1], size=od.shape[0], p=[1 - reroute_pct, reroute_pct])
            if "type" not in od.columns:
                od["type"] = None
            frames.append(od)
        all_od = pd.concat(frames, sort=False, ignore_index=True)
        all_od = all_od.sample(frac=1).reset_index(drop=True)
        print("# agents from file ", all_od.shape[0])
        for row in all_od.itertuples():
            self.agents[row.agent_id] = Agent(row.agent_id, row.origin_nid, row.destin_nid, row.dept_time,
                                              row.veh_len, row.gps_reroute, row.type)

    # ---------------------------------------------------------------------------------------------------------
    # routing in batches (Agent.get_path for many agents: one GPU tree per distinct start node)
    # ---------------------------------------------------------------------------------------------------------
    def route_agents(self, t, agents=None):
        """`agent.get_path(t, g)` + `agent.find_next_link(...)` (run_traffic_simulation.py:107-109) for `agents`
        (default: all), computed as one `sp_route_batch` call.  Returns the agents whose route changed."""
        agents = list(self.agents.values()) if agents is None else list(agents)
        if not agents:
            return []
        # one tree per distinct start node, computed and walked on the GPU: only distances and paths come back
        dist, off, ln, pool = self.g.route_batch([a.current_link_end_nid for a in agents], [a.destin_nid for a in agents])
        changed = []
        for i, a in enumerate(agents):
            d = float(dist[i])
            hops = []
            if d <= 1e8:
                path = pool[off[i]: off[i] + ln[i]].tolist()
                hops = list(zip(path[:-1], path[1:]))
            before, previous_distance = dict(a.route), a.route_distance
            a._adopt(t, d, hops)
            a.find_next_link(node2link_dict=self.node2link_dict)
            if a.route != before:
                changed.append(a)
            # Butte's detour rule (model/dta_meso_butte.py:199-204): the new route is more than `detour_factor`
            # times the old one -> the agent shelters in place (shelter_a2); applied by _reroute_due
            if self.detour_factor and a.status == "enroute" and a.route_distance > previous_distance * self.detour_factor:
                self._detoured.append(a)
        return changed

    def shelter_agents(self, agent_ids, status, t):
        """What the drivers do when one of their rules stops an agent on the road (shelter_a2 after a detour,
        model/dta_meso_butte.py:199-204, 256-263; shelter_park / shelter_a1, run_traffic_simulation.py:113-118):
        the agent takes `status`, leaves its link's queue / run list at once and moves to `agents_stopped`."""
        agent_ids = list(agent_ids)
        if not agent_ids:
            return
        idx = [self._flat.agent_index[i] for i in agent_ids]
        self.sim.stop_agents(idx, STATUS_CODES[status], t)
        for i, aid in zip(idx, agent_ids):
            a = self._agent_objs[i]
            a.status = status
            self._arrive_reported[i] = True
            self.agents_stopped[aid] = (status, t, a.agent_type)
            self.agents.pop(aid, None)

    # ---------------------------------------------------------------------------------------------------------
    # the device switch
    # ---------------------------------------------------------------------------------------------------------
    def use_gpu(self, seed=0, special_nodes=None, reroute_freq=300, weight_freq=100, route_at_start=True,
                route_mode="device", shelter_after=3 * 3600, congest_after=1800, detour_factor=None,
                closure_reroute_freq=None):
        """Move the simulation onto the GPU.  Call once after `add_demand` (and after any t=0 changes to
        `agent.departure_time`).

        route_mode="device" (default): everything the driver's agent loop does stays on the GPU -- the edge-weight
        refresh every `weight_freq` ticks, one reverse shortest-path tree per destination, the staggered reroute
        of agents with t % reroute_freq == agent_id % reroute_freq, loading, arrivals and the give-up rule after
        `shelter_after` seconds on one link, the per-tick `link.congested` count of vehicles more than
        `congest_after` seconds on their link (run_traffic_simulation.py:80-118).  `network.step(t)` is one C call.
        route_mode="host": routes are Python dicts on the agents like in the reference (`agent.route`), recomputed
        in batches of forward trees and pushed to the device; use it when a driver edits routes itself
        (e.g. get_partial_path).  Butte's two extra rules (model/dta_meso_butte.py:199-209) work in both modes:
        `detour_factor=2` shelters an enroute agent whose rerouted trip is more than twice its previous one
        (shelter_a2), `closure_reroute_freq=10` reroutes agents whose next link is closed when
        t % 10 == agent_id % 10.  In device mode they run inside the router (sq_set_reroute_rules), in host mode in
        Python (`route_agents`, `shelter_agents`); either way before the link model of the tick, i.e. before the
        give-up test of the same second (the reference tests give-up first; the two only meet when an agent's
        stagger slot falls on the very second it gives up)."""
        from .flatten import router_arrays
        from .tick import DeviceSim
        if special_nodes is not None:
            self.special_nodes = special_nodes
        if route_mode not in ("device", "host"):
            raise ValueError("route_mode must be 'device' or 'host'")
        self.route_mode = route_mode
        self._pending_status = {}
        self.detour_factor, self.closure_reroute_freq = detour_factor, closure_reroute_freq
        if route_mode == "host":
            if route_at_start:
                self.route_agents(0)
            self._flat = flatten_network(self, self.special_nodes)
            self._sim = DeviceSim(self._flat, seed=seed)
        else:
            # the t == 0 reachability rule (queue_class.py:494-499) needs the distances to every destination once
            dests = sorted({a.destin_nid for a in self.agents.values()})
            slot = {d: k for k, d in enumerate(dests)}
            dist, _ = self.g.sssp_batch(np.array(dests, np.int32), want_parent=False, reverse=True, pinned=False)
            for a in self.agents.values():
                if dist[slot[a.destin_nid], a.origin_nid] > 1e8:
                    a.route, a.status = {}, "shelter_p"
            self._flat = router_arrays(flatten_network(self, self.special_nodes), self)
            self._sim = DeviceSim(self._flat, seed=seed, routes_from_router=True)
            self._sim.attach_router(self.g, self._flat, reroute_freq=reroute_freq, weight_freq=weight_freq)
            if detour_factor or closure_reroute_freq:
                # Butte's two rules inside the device router (sq_set_reroute_rules): no per-tick host work
                self._sim.set_reroute_rules(detour_factor or 0.0, closure_reroute_freq or 0)
        if shelter_after:
            self._sim.set_shelter_after(shelter_after)
        if congest_after:
            self._sim.set_congest_after(congest_after)
        for link in self.links.values():
            link._net = self
        self._seed = seed
        self.reroute_freq = reroute_freq
        self.weight_freq = weight_freq
        self._agent_objs = list(self.agents.values())
        self._link_objs = list(self.links.values())
        self._real_edges = [(i, l) for i, l in enumerate(self._link_objs) if l.link_type != "v"]
        self._end_node_of_link = [l.end_nid for l in self._link_objs]
        self._arrive_reported = np.zeros(len(self._agent_objs), bool)
        return self

    @property
    def sim(self):
        if self._sim is None:
            raise B200LibraryError("network.use_gpu() has not been called")
        return self._sim

    def refresh_edge_weights(self):
        """`link.update_travel_time_by_queue_length(g)` for every link (run_traffic_simulation.py:83-84,
        queue_class.py:452-455) from the device queue lengths: one bulk read, one update per open real link."""
        w = self.sim.link_weights()
        for i, link in self._real_edges:
            if not math.isnan(w[i]):
                self.g.update_edge(link.start_nid, link.end_nid, float(w[i]))

    def _reroute_due(self, t):
        """Agents with t % reroute_freq == agent_id % reroute_freq that are still travelling or waiting to
        depart (run_traffic_simulation.py:95-98, 107)."""
        f, cf = self.reroute_freq, self.closure_reroute_freq if self._closed_idx else None
        if not f and not cf:
            return
        due = [i for i, a in enumerate(self._agent_objs) if f and a.agent_id % f == t % f]
        # reroute upon closure (model/dta_meso_butte.py:205-209): the agent's next link is closed, staggered by id
        near = [i for i, a in enumerate(self._agent_objs) if cf and a.agent_id % cf == t % cf]
        if not due and not near:
            return
        st = self.sim.agent_state()
        live = [i for i in due if st["status"][i] in (0, 1)]
        seen = set(live)
        live += [i for i in near if i not in seen and st["status"][i] in (0, 1) and int(st["nxt"][i]) in self._closed_idx]
        if not live:
            return
        for i in live:
            self._pull_agent_position(self._agent_objs[i], st, i)
        changed = self.route_agents(t, [self._agent_objs[i] for i in live])
        if changed:
            idx, ptr, links = [], [0], []
            limit = len(self._link_objs) + 1
            for a in changed:
                seq = route_to_links(a.route, a.current_link_end_nid, self.node2link_dict, self._flat.link_index, limit)
                idx.append(self._flat.agent_index[a.agent_id])
                links += seq
                ptr.append(len(links))
            self.sim.set_routes(idx, ptr, links)
        if self._detoured:
            gone, self._detoured = self._detoured, []
            self.shelter_agents([a.agent_id for a in gone], "shelter_a2", t)

    def _pull_agent_position(self, a, st, i):
        cur = int(st["cur"][i])
        if cur >= 0:
            link = self._link_objs[cur]
            a.current_link_start_nid, a.current_link_end_nid = link.start_nid, link.end_nid
            a.current_link_enter_time = int(st["enter"][i])
        a.status = STATUS_NAMES[int(st["status"][i])]

    def step(self, t):
        """One simulated second: the edge-weight refresh every `weight_freq` ticks, the staggered reroute
        trigger, then load + link model + node model on the device (sq_step)."""
        sim = self.sim
        self._flush_link_status()
        if self.route_mode == "host":
            if self.weight_freq and t % self.weight_freq == 0 and t > 0:
                self.refresh_edge_weights()
            if t > 0:
                self._reroute_due(t)
        sim.step(t)

    def count_link_pair_flows(self):
        """Start counting vehicles per (in_link, out_link) pair on the device: the second return value of
        Node.run_node_model (model/queue_class.py:377-383) that Butte's node_model accumulates into
        check_traffic_flow_links_dict (model/dta_meso_butte.py:124-131).  Call after use_gpu, before the first step."""
        self.sim.enable_pair_moves()
        return self

    def link_pair_flows(self, pairs=None):
        """{(in_link_id, out_link_id): vehicles moved so far}; `pairs` restricts the result to the given keys (the
        reference's check_traffic_flow_links_dict), with 0 for pairs nobody crossed."""
        il, ol, cnt = self.sim.pair_moves(self._flat)
        ids = self._flat.link_ids
        flows = {(ids[a], ids[b]): int(c) for a, b, c in zip(il, ol, cnt)}
        if pairs is not None:
            return {k: flows.get(k, 0) for k in pairs}
        return flows

    def summary(self):
        """Counts for the drivers' periodic log / t_stats line (model/dta_meso_butte.py:291-348) from one small
        device reduction instead of a full sync_to_host: agents by status, vehicles queued / running / congested /
        on closed links, closed links, node moves."""
        return self.sim.summary()

    def schedule_closures(self, events):
        """Timed closures as a device event table: events = [(tick, link_id, closed)], the timed part of
        model/dta_meso_butte.py:88-115 without a host call per tick.  At the start of tick `tick` the link stops
        (or resumes) taking vehicles and its routing weight becomes 1e8 (or its free-flow time), exactly as if the
        driver had called close_link_to_newcomers / open_link_to_newcomers and set `status` before `step(tick)`.
        Device route mode only (the host router reads the graph the driver edits itself)."""
        if self.route_mode != "device":
            raise B200LibraryError("schedule_closures needs route_mode='device'")
        ev = sorted(((int(t), lid, bool(c)) for t, lid, c in events), key=lambda e: e[0])
        self.sim.set_closure_schedule([e[0] for e in ev], [self._flat.link_index[e[1]] for e in ev],
                                      [1 if e[2] else 0 for e in ev])
        self._closure_events = ev

    def _flush_link_status(self):
        if self._pending_status:
            idx = [self._flat.link_index[lid] for lid in self._pending_status]
            sim = self.sim
            sim.set_link_closed(idx, [1 if c else 0 for c in self._pending_status.values()])
            for i, c in zip(idx, self._pending_status.values()):
                (self._closed_idx.add if c else self._closed_idx.discard)(i)
            self._pending_status.clear()

    def run(self, t0, n):
        """n consecutive ticks.  In device route mode this is one C call (CUDA-graph replay between the routing
        refreshes); in host route mode the host reroutes between ticks."""
        if self.route_mode == "device":
            self._flush_link_status()
            self.sim.run(t0, n)
        else:
            for t in range(t0, t0 + n):
                self.step(t)

    def sync_to_host(self, t=None):
        """Refresh the object fields drivers read: link vehicle lists / counters / capacities,
        agent status and position, node_move; arrived agents move to `agents_stopped` as the reference's agent
        loop does (run_traffic_simulation.py:95-98, 119-120)."""
        sim = self.sim
        ls, ag, nd, sh = sim.link_state(), sim.agent_state(), sim.node_state(), sim.shelter_state()
        cong = sim.link_congested()
        ptr, veh = sim.link_vehicles()
        ids = self._flat.agent_ids
        for i, link in enumerate(self._link_objs):
            on = [ids[v] for v in veh[ptr[i]: ptr[i + 1]]]
            q = int(ls["q"][i])
            link.queue_vehicles, link.run_vehicles = on[:q], on[q:]
            link.total_vehicles_left = int(ls["left"][i])
            link.congested = int(cong[i])
            link.remaining_inflow_capacity = float(ls["in_c"][i])
            link.remaining_outflow_capacity = float(ls["out_c"][i])
            link.remaining_storage_capacity = float(ls["storage_c"][i])
        for i, a in enumerate(self._agent_objs):
            self._pull_agent_position(a, ag, i)
            cur, nxt = int(ag["cur"][i]), int(ag["nxt"][i])
            a.this_link = self._flat.link_ids[cur] if cur >= 0 else None
            a.next_link = self._flat.link_ids[nxt] if nxt >= 0 else None
            a.next_link_end_nid = self._end_node_of_link[nxt] if nxt >= 0 else None
            if a.status in ("arrive", "shelter_a1", "shelter_a2") and not self._arrive_reported[i]:
                self._arrive_reported[i] = True
                # the reference's agent loop notices an arrival one tick later, a give-up in the same tick
                when = int(sh["stop_tick"][i]) + (1 if a.status == "arrive" else 0)
                self.agents_stopped[a.agent_id] = (a.status, when, a.agent_type)
                self.agents.pop(a.agent_id, None)
        for k, nid in enumerate(self._flat.node_ids):
            self.nodes[nid].node_move = int(nd["node_move"][k])
            self.nodes[nid].shelter_counts = 0
        for i, link in enumerate(self._link_objs):
            if sh["link_shelter"][i]:
                self.nodes[link.end_nid].shelter_counts += int(sh["link_shelter"][i])
        return self
