/*
 * sq_tick.h -- C ABI of the B200 spatial-queue tick (HP-1, SURVEY.md section 8 rows A0-A8, seam B2).
 *
 * One call, sq_step(sim, t), replaces what the reference drivers do once per simulated second
 * (projects/bolinas_civic/run_traffic_simulation.py:111, 52-69):
 *     for agent in network.agents.values(): agent.load_vehicle(t, ...)          queue_class.py:587-596
 *     for link  in network.links.values():  link.run_link_model(t, ...)         queue_class.py:445-455
 *     for node  in {end node of links with a queue}: node.run_node_model(t,...) queue_class.py:226-394
 * State lives on the GPU between calls (SoA arrays, per-link vehicle rings, per-node turn tables);
 * the getters below copy it back for the object mirror (spatial_queue_b200/queue_class.py) at
 * whatever cadence the driver reads it.
 *
 * The network description is the flat form of the reference's objects, in the reference's own
 * iteration orders (spatial_queue_b200/flatten.py).  Indices: real nodes 0..n_nodes-1; links
 * 0..n_links-1 in network.links order (virtual source links first); agents in network.agents order.
 *
 * Randomness: the reference's shuffle / random.choices draws are replaced by the counter-based
 * contract of spatial_queue_b200/rng.py, keyed on (seed, t, node_key, slot).
 *
 * Plain pointers and sizes only; host pointers unless a name ends in _dev.  No CPU fallback.
 */
#ifndef SQ_TICK_H_
#define SQ_TICK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SqNetDesc {
  int32_t n_nodes, n_links, n_agents;
  int32_t reserved0;               /* flags: bit 0 = agents may come without routes (sq_attach_router follows) */
  uint64_t seed;
  /* ---- links (queue_class.py:396-442) ---- */
  const int32_t* link_start_node;  /* real node index, -1 for the virtual start of a source link     */
  const int32_t* link_end_node;
  const uint8_t* link_is_virtual;  /* link_type == 'v' (queue_class.py:224)                           */
  const int32_t* link_lanes_floor; /* int(floor(lanes)) (:246, :279)                                  */
  const double* link_cap_per_s;    /* capacity / 3600 as Python computes it (:454)                    */
  const double* link_storage_cap;  /* max(18, length * lanes) (:416)                                  */
  const double* link_fft;          /* free-flow time, s (:54)                                         */
  const uint8_t* link_closed;      /* status in {closed, burning_closed} (:455)                       */
  const int32_t* link_reverse;     /* node2link[(end, start)] or -1 (:272)                            */
  const double* link_capacity;     /* veh/h, for the edge-weight model (:465)                         */
  const double* link_maxmph;       /* (:411)                                                          */
  /* ---- nodes (queue_class.py:204-218, 158-176) ---- */
  const int64_t* node_key;         /* integer node id, RNG key                                        */
  const uint8_t* node_type;        /* 0 plain, 1 signal, 2 priority (:233-238)                        */
  const int32_t* node_in_ptr;      /* [n_nodes+1] CSR over incoming_links order                       */
  const int32_t* node_in_links;
  const uint8_t* node_in_phase;    /* per in-link: bit0 in phase_1, bit1 in phase_2 (signal nodes)    */
  const int32_t* node_in_prio;     /* per in-link: rank in the priority list or -1 (priority nodes)   */
  const int32_t* node_out_ptr;     /* [n_nodes+1] CSR over outgoing_links order                       */
  const int32_t* node_out_links;
  const int64_t* node_turn_ptr;    /* [n_nodes+1] offset of the node's indeg x outdeg angle table      */
  const double* turn_angle;        /* degrees, row = in-link, col = out-link (:169-176)               */
  /* ---- agents (queue_class.py:478-499) ---- */
  const int64_t* agent_key;
  const int32_t* agent_origin;
  const int32_t* agent_destin;
  const int32_t* agent_dept;
  const int32_t* agent_vlen;       /* 8 or 18 (:191)                                                  */
  const uint8_t* agent_status;     /* 0 unloaded ... (flatten.STATUS_CODES); 3 = shelter_p = inactive */
  const int64_t* route_ptr;        /* [n_agents+1]                                                    */
  const int32_t* route_links;      /* link indices from the origin node to the destination           */
} SqNetDesc;

typedef struct SqSim SqSim;

/* Agent status codes (spatial_queue_b200/flatten.py STATUS_CODES). */
enum { SQ_UNLOADED = 0, SQ_ENROUTE = 1, SQ_ARRIVE = 2, SQ_SHELTER_P = 3, SQ_SHELTER_A1 = 4, SQ_SHELTER_A2 = 5 };

/* Build device state from a network that has not ticked yet.  NULL on failure (sq_last_error). */
SqSim* sq_create(const SqNetDesc* desc);
void sq_destroy(SqSim* sim);

/* One simulated second: load + link model + node model for tick t.  Asynchronous; errors detected on
 * the device (a queued vehicle with no next link away from its destination, queue_class.py:312-314;
 * a turn that is not in the node's table, :257-262) surface from sq_sync / the getters as SQ_ERR_STATE. */
int sq_step(SqSim* sim, int32_t t);
/* n consecutive ticks t0 .. t0+n-1 (replayed as CUDA graphs of tick pairs). */
int sq_run(SqSim* sim, int32_t t0, int32_t n);
/* Run all later work of this sim on `stream` (a cudaStream_t passed as void*, e.g. torch's current stream)
 * instead of the sim's own stream, so a caller can bracket it with its own events. */
int sq_set_stream(SqSim* sim, void* stream);
/* Wait for the device and report any device-side error. */
int sq_sync(SqSim* sim);

/* Replace the routes of n agents (Agent.get_path + find_next_link, queue_class.py:501-522, 598-608):
 * ptr[n+1] offsets into links[]; the new route starts at the end node of the agent's current link. */
int sq_set_routes(SqSim* sim, int32_t n, const int32_t* agent_idx, const int64_t* ptr, const int32_t* links);
/* Link status changes (model/dta_meso_butte.py:77-81, 89-101, 113): closed != 0 zeroes the inflow
 * capacity in the link model (queue_class.py:455). */
int sq_set_link_closed(SqSim* sim, int32_t n, const int32_t* link_idx, const uint8_t* closed);
/* Closure schedule as a device event table (the timed closures and reopenings of
 * model/dta_meso_butte.py:88-115): event i sets link_idx[i] closed (closed[i] != 0) or open at the start of tick
 * tick[i], before that tick's link phase -- the same moment a driver would call Link.close_link_to_newcomers /
 * open_link_to_newcomers and set Link.status before its link_model(t).  With a router attached the routing graph
 * gets 1e8 / the free-flow time on the link's edge (queue_class.py:467-474) and the next-link tables are rebuilt
 * at once, all on the device: sq_run keeps replaying CUDA graphs between events.  Replaces any earlier schedule;
 * events earlier than the next tick run are applied at that tick. */
int sq_set_closure_schedule(SqSim* sim, int32_t n, const int32_t* tick, const int32_t* link_idx, const uint8_t* closed);

/* ---- device-side rerouting (Agent.get_path for every agent, queue_class.py:501-522; the drivers' reroute
 * trigger and edge-weight refresh, run_traffic_simulation.py:81-84, 107-109) --------------------------------
 * With a router attached, sq_step / sq_run do by themselves what the driver's agent loop does:
 *   every weight_freq ticks: link weights from the queue lengths (queue_class.py:463-466) -> the routing graph
 *                            (sp_set_weights_device), one reverse shortest-path tree per destination
 *                            (sp_sssp_batch_device_ex) -> next-link tables;
 *   every tick t > 0:        agents with t % reroute_freq == agent_key % reroute_freq adopt the newest tables from
 *                            the end node of their current link unless the destination is out of reach (> 1e8).
 * Attaching routes every agent once on the weights of the empty network (the t == 0 rule).  Agents whose
 * destination is unreachable at t == 0 must be passed to sq_create as SQ_SHELTER_P (queue_class.py:504-510). */
typedef struct SqRouterDesc {
  void* graph;                 /* Graph* of include/sq_sp.h holding the real links as edges                       */
  int32_t n_graph_edges;       /* edges in creategraph order                                                      */
  int32_t n_dest;
  const int32_t* dest_nodes;   /* [n_dest] node indices                                                           */
  const int32_t* agent_dslot;  /* [n_agents] index into dest_nodes                                                */
  const int64_t* agent_key;    /* [n_agents] agent ids (reroute stagger)                                          */
  const int32_t* node_vertex;  /* [n_nodes] graph vertex id of each node                                          */
  const int32_t* link_edge;    /* [n_links] graph edge that receives the link's weight, -1 = none                 */
  int32_t n_pairs;             /* node2link_dict restricted to real start nodes (queue_class.py:149-156)          */
  const int32_t* pair_start; const int32_t* pair_end; const int32_t* pair_link;
  int32_t reroute_freq, weight_freq;
} SqRouterDesc;
int sq_attach_router(SqSim* sim, const SqRouterDesc* router);

/* Butte's additions to the agent loop (model/dta_meso_butte.py:193-209, 256-263), inside the device router:
 *   detour_factor > 0:        whenever an agent is rerouted, its new route distance is compared with the one it had
 *                             (Agent.route_distance, 5e7 before the first route); an ENROUTE agent whose trip grew by
 *                             more than the factor (the reference uses 2) gets status SQ_SHELTER_A2 and leaves its
 *                             link's queue / run list before the link model of the same tick;
 *   closure_reroute_freq > 0: an agent whose next link is closed (sq_set_link_closed / closure schedule) reroutes when
 *                             t % closure_reroute_freq == agent_key % closure_reroute_freq (the reference uses 10),
 *                             after its regular reroute if both fall on the same tick.
 * The reroutes then run as a launch of their own before the link phase (three kernels per tick instead of two).
 * Call once, after sq_attach_router and before the first tick. */
int sq_set_reroute_rules(SqSim* sim, double detour_factor, int32_t closure_reroute_freq);

/* Agent-loop rule of the drivers (run_traffic_simulation.py:100-104, 113-118): a vehicle that has spent more than
 * `seconds` on one real link gives up (status SQ_SHELTER_A1) and leaves the link before the link model of that tick.
 * 0 (default) switches the rule off. */
int sq_set_shelter_after(SqSim* sim, int32_t seconds);
/* A driver rule took agents off the road (Butte's shelter_a2 detour rule, model/dta_meso_butte.py:199-263; the
 * "remove passively sheltered vehicles immediately" step of both drivers, run_traffic_simulation.py:113-118): the
 * listed ENROUTE agents get `status` (SQ_SHELTER_P or above; 5 = shelter_a2 in spatial_queue_b200/flatten.py) and
 * stop tick `t`, and leave the queue / run list of their link at once, the other vehicles keeping their order.
 * Call between ticks.  An agent that is not enroute makes the call fail (SQ_ERR_STATE). */
int sq_stop_agents(SqSim* sim, int32_t n, const int32_t* agent_idx, int32_t status, int32_t t);

/* The congestion counter of the same loop (run_traffic_simulation.py:100-102; model/dta_meso_butte.py:183-185 uses
 * max(1800, 10 fft)): link.congested is reset every tick (run_traffic_simulation.py:80-82) and then counts the vehicles that have
 * been on the (real) link for more than `seconds`.  0 (default) switches it off; sq_get_link_congested reads the
 * per-link counts of the last tick run. */
int sq_set_congest_after(SqSim* sim, int32_t seconds);
int sq_get_link_congested(SqSim* sim, int64_t* congested);
/* Output cadence in one small read-back (the t_stats / link_stats lines of
 * model/dta_meso_butte.py:291-348): out16 = [0..7] agents by status code (SQ_UNLOADED ... SQ_SHELTER_A1),
 * [8] queued vehicles, [9] running vehicles, [10] congested vehicles (last tick), [11] closed links,
 * [12] sum of node_move, [13] sum of total_vehicles_left over real links, [14] real links with a vehicle on them,
 * [15] vehicles on closed links (Butte's in_fire_cnts for burning links). */
int sq_get_summary(SqSim* sim, int64_t* out16);
/* The second return value of Node.run_node_model (model/queue_class.py:287-383: node_move_link_pairs, one (in_link,
 * out_link) pair per vehicle moved) as counters: after sq_enable_pair_moves every move from in-link i to out-link j of
 * a node adds one to entry node_turn_ptr[node] + i_local * outdeg + j_local of a table shaped like the turn-angle table
 * (SqNetDesc.turn_angle); sq_get_pair_moves reads it back (out holds node_turn_ptr[n_nodes] entries).  Butte's
 * check_traffic_flow_links_dict (model/dta_meso_butte.py:124-131) is a lookup in that table.  Arrivals, which the
 * reference records as (in_link, None), are total_vehicles_left[in_link] minus the row sum. */
int sq_enable_pair_moves(SqSim* sim);
int sq_get_pair_moves(SqSim* sim, int64_t* out);
/* Measurement aid, no reference counterpart: record %globaltimer (ns) at entry and exit of every warp of the two phase
 * kernels for ticks [t0, t0 + n_ticks).  sq_get_timeline returns the window and the row stride (warps), then -- when
 * out is not null -- n_ticks x 2 (link phase, node phase) x stride pairs (entry, exit); unused slots are (0, 0). */
int sq_enable_timeline(SqSim* sim, int32_t t0, int32_t n_ticks);
int sq_get_timeline(SqSim* sim, int32_t* n_ticks, int32_t* stride, uint64_t* out);
/* Test hook, host only (no GPU needed), no reference counterpart: the static node records the node kernel reads before
 * its dependent chain -- per node 16 ints: the first eight in-links (-1 beyond the in-degree), then per in-link slot i the
 * packed order in which the other slots are tried as the opposite approach when i is the primary link (bits 0..2 count,
 * 3 bits per entry from bit 3; bit 30 = link i has no reverse link, bit 31 = ties / duplicate ranks: the kernel walks
 * the angle table as queue_class.py:267-293 does; bit 29 of slot 0 = more than eight in-links).
 * tests/test_node_records.py checks the lists against that walk for every candidate set. */
int sq_build_node_records(const SqNetDesc* desc, int32_t* out);
/* Per link: vehicles that sheltered on it (sum over a node's in-links = node.shelter_counts); per agent: the tick
 * at which it arrived or sheltered (-1 otherwise). */
int sq_get_shelter_state(SqSim* sim, int32_t* link_shelter, int32_t* agent_stop_tick);

/* ---- state read-back (any pointer may be NULL) ---- */
int sq_get_link_state(SqSim* sim, int32_t* queue_len, int32_t* run_len, int32_t* total_left, double* inflow_c,
                      double* outflow_c, double* storage_c);
/* Vehicles on every link, queue first then run list, in list order: ptr[n_links+1], agents[ptr[n_links]]
 * (agents must hold n_agents entries). */
int sq_get_link_vehicles(SqSim* sim, int64_t* ptr, int32_t* agents);
int sq_get_agent_state(SqSim* sim, uint8_t* status, int32_t* cur_link, int32_t* next_link, int32_t* enter_time,
                       int32_t* arrive_tick);
int sq_get_node_state(SqSim* sim, int64_t* node_move);
/* Edge weights for routing, one per link (queue_class.py:463-466): len(queue)/(capacity+0.1)*3600*25/maxmph + fft
 * for open real links, else the current weight is left untouched (returned as NaN). */
int sq_get_link_weights(SqSim* sim, double* weights);
/* Run counters: [0] ticks, [1] sum over ticks of vehicles on links after the tick (agent-steps),
 * [2] moves (transfers + arrivals), [3] active nodes, [4] arrivals, [5] kernel launches of this sim. */
int sq_get_counters(SqSim* sim, int64_t* out6);

const char* sq_last_error(void);
long long sq_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* SQ_TICK_H_ */
