"""Network / Node / Link / Agent objects -> flat arrays (the boundary of SURVEY.md 8 row A0).

Works on any object graph with the reference's attribute names (model/queue_class.py:24-32, 204-218, 396-442,
478-499), i.e. both the reference's own `Network` and this package's mirror.  Every order the tick's results
depend on is taken from the objects as the reference iterates them:

  links      network.links dict order (virtual links in node-CSV order, then real links; queue_class.py:137-148)
  in-links   node.incoming_links dict order (queue_class.py:159-161), out-links node.outgoing_links
  agents     network.agents dict order (the shuffled OD rows; queue_class.py:198-202)
"""
import numpy as np

STATUS_CODES = {"unloaded": 0, "enroute": 1, "arrive": 2, "shelter_p": 3, "shelter_a1": 4, "shelter_a2": 5,
                "shelter_park": 6}
STATUS_NAMES = {v: k for k, v in STATUS_CODES.items()}
CLOSED_STATUSES = ("closed", "burning_closed")  # queue_class.py:455


class FlatNetwork:
    """Plain container of numpy arrays; field docs in include/sq_tick.h (SqNetDesc)."""

    def __init__(self):
        self.link_ids = []    # python link ids in index order
        self.node_ids = []    # python ids of the real nodes in index order
        self.agent_ids = []   # python agent ids in index order

    def as_dict(self):
        return {k: v for k, v in self.__dict__.items() if isinstance(v, np.ndarray)}

    @classmethod
    def from_dict(cls, d):
        f = cls()
        for k, v in d.items():
            setattr(f, k, np.asarray(v))
        return f


def route_to_links(route, start_nid, node2link_dict, link_index, limit):
    """Follow a reference route dict {node: next node} from `start_nid` and return the link-index sequence."""
    seq = []
    n = start_nid
    while n in route and len(seq) <= limit:
        nxt = route[n]
        seq.append(link_index[node2link_dict[(n, nxt)]])
        n = nxt
    return seq


def flatten_network(network, special_nodes=None):
    special_nodes = special_nodes or {}
    f = FlatNetwork()
    real_nodes = [(nid, n) for nid, n in network.nodes.items() if getattr(n, "node_type", "real") != "v"]
    f.node_ids = [nid for nid, _ in real_nodes]
    node_index = {nid: i for i, nid in enumerate(f.node_ids)}
    f.link_ids = list(network.links.keys())
    link_index = {lid: i for i, lid in enumerate(f.link_ids)}
    f.node_index, f.link_index = node_index, link_index
    N, L = len(real_nodes), len(f.link_ids)

    f.node_key = np.array([int(nid) for nid in f.node_ids], dtype=np.int64)
    f.link_start_node = np.full(L, -1, np.int32)
    f.link_end_node = np.zeros(L, np.int32)
    f.link_is_virtual = np.zeros(L, np.uint8)
    f.link_lanes_floor = np.zeros(L, np.int32)
    f.link_cap_per_s = np.zeros(L, np.float64)
    f.link_storage_cap = np.zeros(L, np.float64)
    f.link_fft = np.zeros(L, np.float64)
    f.link_closed = np.zeros(L, np.uint8)
    f.link_reverse = np.full(L, -1, np.int32)
    # edge-weight model inputs (queue_class.py:463-466): capacity, maxmph
    f.link_capacity = np.zeros(L, np.float64)
    f.link_maxmph = np.zeros(L, np.float64)
    for i, (lid, link) in enumerate(network.links.items()):
        f.link_start_node[i] = node_index.get(link.start_nid, -1)
        f.link_end_node[i] = node_index[link.end_nid]
        f.link_is_virtual[i] = 1 if link.link_type == "v" else 0
        f.link_lanes_floor[i] = int(np.floor(link.lanes))
        f.link_cap_per_s[i] = link.capacity / 3600
        f.link_storage_cap[i] = float(link.storage_capacity)
        f.link_fft[i] = float(link.fft)
        f.link_closed[i] = 1 if link.status in CLOSED_STATUSES else 0
        rev = network.node2link_dict.get((link.end_nid, link.start_nid))
        f.link_reverse[i] = link_index[rev] if rev is not None else -1
        f.link_capacity[i] = float(link.capacity)
        f.link_maxmph[i] = float(link.maxmph)
        if link.queue_vehicles or link.run_vehicles:
            raise ValueError("flatten_network expects a network before its first tick (vehicle lists not empty)")

    in_ptr, in_links, out_ptr, out_links, turn_ptr, turn = [0], [], [0], [], [0], []
    in_phase, in_prio = [], []
    f.node_type = np.zeros(N, np.uint8)
    sig = special_nodes.get("signal", {}) or {}
    pri = special_nodes.get("priority", {}) or {}
    for i, (nid, node) in enumerate(real_nodes):
        ins = list(node.incoming_links.keys())
        outs = list(node.outgoing_links)
        key = str(nid)
        if key in sig:
            f.node_type[i] = 1
        elif key in pri:
            f.node_type[i] = 2
        for il in ins:
            in_links.append(link_index[il])
            ph = 0
            if key in sig:
                ph = (1 if il in sig[key]["phase_1"] else 0) | (2 if il in sig[key]["phase_2"] else 0)
            in_phase.append(ph)
            in_prio.append(pri[key].index(il) if (key in pri and il in pri[key]) else -1)
            for ol in outs:
                turn.append(float(node.incoming_links[il][ol]))
        out_links += [link_index[ol] for ol in outs]
        in_ptr.append(len(in_links))
        out_ptr.append(len(out_links))
        turn_ptr.append(len(turn))
    f.node_in_ptr = np.array(in_ptr, np.int32)
    f.node_in_links = np.array(in_links, np.int32)
    f.node_in_phase = np.array(in_phase, np.uint8)
    f.node_in_prio = np.array(in_prio, np.int32)
    f.node_out_ptr = np.array(out_ptr, np.int32)
    f.node_out_links = np.array(out_links, np.int32)
    f.node_turn_ptr = np.array(turn_ptr, np.int64)
    f.turn_angle = np.array(turn, np.float64)

    # the (start node, end node) -> link choice the agents use (queue_class.py:149-156), as a per-node CSR
    # over outgoing neighbours; the device router turns predecessor trees into link ids with it
    f.pair_start = np.array([node_index[s] for (s, e) in network.node2link_dict if s in node_index], np.int32)
    f.pair_end = np.array([node_index[e] for (s, e) in network.node2link_dict if s in node_index], np.int32)
    f.pair_link = np.array([link_index[l] for (s, e), l in network.node2link_dict.items() if s in node_index], np.int32)

    flatten_agents(f, network)
    return f


def flatten_agents(f, network):
    agents = list(network.agents.items())
    A = len(agents)
    f.agent_ids = [aid for aid, _ in agents]
    f.agent_index = {aid: i for i, aid in enumerate(f.agent_ids)}
    f.agent_key = np.array([int(aid) for aid in f.agent_ids], np.int64)
    f.agent_origin = np.zeros(A, np.int32)
    f.agent_destin = np.zeros(A, np.int32)
    f.agent_dept = np.zeros(A, np.int32)
    f.agent_vlen = np.zeros(A, np.int32)
    f.agent_status = np.zeros(A, np.uint8)
    ptr, links = [0], []
    limit = len(f.link_ids) + 1
    for i, (aid, a) in enumerate(agents):
        f.agent_origin[i] = f.node_index[a.origin_nid]
        f.agent_destin[i] = f.node_index[a.destin_nid]
        f.agent_dept[i] = int(a.departure_time)
        f.agent_vlen[i] = int(a.vehicle_length)
        f.agent_status[i] = STATUS_CODES[a.status]
        if a.status not in ("unloaded", "shelter_p"):
            raise ValueError("flatten_agents expects agents before their first tick")
        links += route_to_links(a.route, a.current_link_end_nid, network.node2link_dict, f.link_index, limit)
        ptr.append(len(links))
    f.route_ptr = np.array(ptr, np.int64)
    f.route_links = np.array(links, np.int32)
    return f


def router_arrays(f, network):
    """Arrays sq_attach_router needs (include/sq_tick.h SqRouterDesc), from a flattened network whose routing graph
    `network.g` was built by the loader (interface.from_dataframe on the real links in links order,
    queue_class.py:123): destinations, per-agent destination slot, node -> graph vertex, link -> graph edge.

    link_edge follows what `link.update_travel_time_by_queue_length(g)` does to parallel links: every link of a
    duplicated (start, end) pair calls update_edge(start, end, w), which writes the LAST inserted edge of the pair
    (graph.cc:13, 34-36), so after the loop over links only the last link's weight survives, on the last edge."""
    f.node_vertex = f.node_key.astype(np.int32)
    n_virtual = int(f.link_is_virtual.sum())
    edge = np.full(len(f.link_ids), -1, np.int32)
    last_of_pair = {}
    for i, lid in enumerate(f.link_ids):
        link = network.links[lid]
        if link.link_type == "v":
            continue
        last_of_pair[(link.start_nid, link.end_nid)] = i
    for i in last_of_pair.values():
        edge[i] = i - n_virtual
    f.link_edge = edge
    dests = sorted(set(int(d) for d in f.agent_destin))
    slot = {d: k for k, d in enumerate(dests)}
    f.dest_nodes = np.array(dests, np.int32)
    f.agent_dslot = np.array([slot[int(d)] for d in f.agent_destin], np.int32)
    return f


def router_arrays_flat(f):
    """router_arrays from the flat arrays alone (golden files, synthetic networks): the routing graph is taken to
    hold the real links in link order (edge k = k-th real link), as the loader builds it."""
    f.node_vertex = np.asarray(f.node_key).astype(np.int32)
    virt = np.asarray(f.link_is_virtual).astype(bool)
    real = np.nonzero(~virt)[0]
    edge = np.full(len(virt), -1, np.int32)
    last = {}
    for k, i in enumerate(real):
        last[(int(f.link_start_node[i]), int(f.link_end_node[i]))] = (i, k)
    for i, k in last.values():
        edge[i] = k
    f.link_edge = edge
    dests = sorted(set(int(d) for d in f.agent_destin))
    slot = {d: k for k, d in enumerate(dests)}
    f.dest_nodes = np.array(dests, np.int32)
    f.agent_dslot = np.array([slot[int(d)] for d in f.agent_destin], np.int32)
    return f


def routing_graph_flat(f):
    """The loader's routing graph rebuilt from flat arrays: real links in order, vertices = node keys, weight = fft."""
    from . import sp_interface
    real = np.nonzero(~np.asarray(f.link_is_virtual).astype(bool))[0]
    key = np.asarray(f.node_key)
    return sp_interface.from_arrays(key[f.link_start_node[real]], key[f.link_end_node[real]], f.link_fft[real],
                                    max_vertex_id=int(key.max()))
