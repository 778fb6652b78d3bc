/*
 * mazetd_b200.h -- C ABI of libmazetd_b200.so: the B200-native (sm_100a) replacement for mazetd's
 * flow-field hot path. Plain pointers and sizes only; no C++ or torch types cross this boundary.
 *
 * Each entry point names the reference interface it replaces (paths relative to the mazetd tree).
 * The reference has no plugin registry: the seam is link-time substitution of the bodies of
 * sources/map/paths.cpp, sources/map/grid.cpp and the elevation part of sources/map/procedural.cpp
 * (SURVEY.md 8b). mazetd_b200/host/ holds those replacement bodies (thin adapters onto this ABI);
 * INTEGRATION.md shows how a maintainer links them.
 *
 * Conventions
 *   - tile index idx = y * W + x (sources/map/grid.cpp:105); MTD_M = 0xFFFFFFFF is cage's `m`.
 *   - distances are uint32 "24.8" fixed point with 255 units per world unit (sources/common.h:15-29).
 *   - flags are mazetd TileFlags bits (sources/grid.h:15-28); pathing reads Invalid|Building only
 *     (sources/map/paths.cpp:65).
 *   - every call returns 0 on success, a negative MTD_ERR_* otherwise; mtd_last_error() has the text.
 *     No C++ exception crosses the boundary (the adapter turns codes into cage::Exception).
 *   - "host" entry points take caller-owned HOST pointers and are synchronous (results are complete on
 *     return, like the reference's update()). "_dev" entry points take DEVICE pointers, enqueue on the
 *     context's stream and do not synchronise: they are what bench.py times for the HBM-resident metric.
 *   - there is no CPU fallback: without a CUDA device mtd_create() fails with MTD_ERR_CUDA.
 */
#ifndef MAZETD_B200_H
#define MAZETD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTD_M 0xFFFFFFFFu
#define MTD_MAX_WAYPOINTS 32 /* visited bits are a uint32 (sources/map/paths.cpp:147) */

enum
{
	MTD_OK = 0,
	MTD_ERR_ARG = -1,      /* bad argument (null pointer, size, tile out of range) */
	MTD_ERR_CUDA = -2,     /* CUDA runtime error, no device, or kernel fault */
	MTD_ERR_STATE = -3,    /* call order (e.g. update before grid upload) */
	MTD_ERR_CAPACITY = -4, /* map does not fit the selected path / buffer too small */
	MTD_ERR_OVERFLOW = -5, /* a distance reached the uint32 guard band (reference would wrap) */
	MTD_ERR_NOMEM = -6
};

typedef struct mtd_ctx mtd_ctx;

/* ---- lifecycle ------------------------------------------------------------------------------- */
/* device: CUDA ordinal. stream: a cudaStream_t to enqueue on (e.g. torch's current stream), or NULL to let
 * the context create its own non-blocking stream. */
int mtd_create(mtd_ctx **out, int device, void *stream);
void mtd_destroy(mtd_ctx *ctx);
const char *mtd_last_error(const mtd_ctx *ctx);
int mtd_version(void);
/* blocks until everything enqueued by this context has finished; returns deferred kernel status */
int mtd_synchronize(mtd_ctx *ctx);
/* number of kernels this context has launched since creation (bench.py's gpu_launches) */
uint64_t mtd_launch_count(const mtd_ctx *ctx);

/* ---- kernel timing probe (bench.py's roofline: per-kernel device time measured live with CUDA events on the
 *      context's stream; replaces the reference's ProfilingScope, externals/cage/sources/include/cage-core/profiling.h:37-51) */
enum
{
	MTD_K_WAYPOINTS_SMALL = 0, /* fused Waypoints::update of shared-memory sized maps */
	MTD_K_MAPGEN_GRID = 1,     /* newProcedural + newGrid + spawn picking */
	MTD_K_BIG_RELAX = 2,       /* tiled relaxation of a large grid (all rounds of one Directions::update) */
	MTD_K_BIG_DIRECTIONS = 3,  /* next-hop pass of a large grid */
	MTD_K_MONSTERS_TICK = 4,
	MTD_K_BIG_PREPARE = 5,     /* edge weights + blocking mask of a large grid */
	MTD_K_COUNT = 6
};
int mtd_profile_enable(mtd_ctx *ctx, int on);
/* synchronises the stream; returns the summed device time and number of bracketed regions of `kind` since the last read */
int mtd_profile_read(mtd_ctx *ctx, int kind, double *total_ms, uint64_t *regions);

/* ---- test hooks (no reference counterpart): force rarely taken kernel paths so the parity suite covers them ------ */
enum
{
	MTD_DEBUG_QSEG_CAP = 1,    /* > 0 caps the per-warp frontier queue segment of the one-CTA path: small values force the queue-less dense rounds */
	MTD_DEBUG_GROUP_WARPS = 2, /* warps cooperating on one distance layer in the one-CTA path: 1, 2 or 4 */
	MTD_DEBUG_WALKER = 3,      /* who walks the full paths in the one-CTA path: 0 auto (walker warp when a CTA owns several maps), 1 workers, 2 walker warp */
	MTD_DEBUG_BIG_ASYNC = 5,   /* large grids: 1 (default) = dataflow scheduler without global rounds (k_tile_async), 0 = round-based kernel (k_tile_rounds). Same result */
	MTD_DEBUG_ASYNC_NB_RULE = 8, /* dataflow scheduler: value >= 0: a tile does not start while a neighbour tile is in flight or waits with a key smaller by more than `value`; < 0: off. Performance only */
	MTD_DEBUG_ASYNC_DEPTH = 9, /* dataflow scheduler: mailboxes per worker (1, or 2 = the next item is handed out and prefetched early). Performance only */
	MTD_DEBUG_INCREMENTAL = 10, /* one-CTA path: 1 (default) = after ONE blocking flip the resident layers are repaired (subtree cut out and re-relaxed / improvement propagated) and the trial reuses layer 0; 0 = always from scratch. Same result */
	MTD_DEBUG_FUSED_MAPGEN = 11, /* mtd_mapgen_batch with waypoints: 1 (default) = the generator runs inside the waypoints kernel (the next map's elevations are computed while the current map is relaxed), 0 = generator kernel, then waypoints kernel. Same result */
	MTD_DEBUG_GEN_WARPS = 12,  /* fused generator: how many of the warps without a layer compute the next map's elevations during the relaxation (0 = all warps, after it). Performance only */
	MTD_DEBUG_ASYNC_WORKERS = 13, /* dataflow scheduler: worker warps per SM that are handed items (1..16). Performance only */
	MTD_DEBUG_ASYNC_LATE_LOOKS = 14, /* dataflow scheduler: how often a visit re-reads its apron from global memory after converging and goes on if a neighbour lowered it meanwhile (0..4, default 1). Performance only */
	MTD_DEBUG_ASYNC_CONT = 15, /* dataflow scheduler: 1 (default) = a revisit of a tile goes to the worker that visited it last (queued behind the running visit, or to that worker if idle) and continues on the frame it left instead of staging it again. Same result */
	MTD_DEBUG_BIG_TMA = 7,     /* dataflow scheduler: 1 (default) = tile frames staged by 2-D bulk tensor copies (TMA), 0 = cp.async rows. Same result */
	MTD_DEBUG_ASYNC_DELTA = 6, /* priority window (label units) of the dataflow scheduler. Performance only */
	MTD_DEBUG_TILE_DELTA = 4   /* large grids: width (label units) of the window of tile offers visited per round; 2^32-1 = visit everything (no priority). Performance only: any value gives the same result */
};
int mtd_debug_set(mtd_ctx *ctx, int key, int64_t value);
/* Statistics of the last large-grid (tiled) update: out4 = { global rounds (0 with the dataflow scheduler), tile visits,
 * scheduler polls, bit 0: the dataflow scheduler ran, bit 1: frames were staged by TMA }. While the context holds a grid of the
 * one-CTA path: { incremental updates, incremental trials since mtd_create, 0, 0 }. Diagnostics for bench.py; no reference counterpart. */
int mtd_big_stats(mtd_ctx *ctx, uint32_t *out4);

/* ---- Grid (replaces the data side of struct Grid, sources/grid.h:38-53) ------------------------ */
/* Upload a grid; builds the device-side derived data (packed edge weights = Grid::neighborDistance for the
 * right/down edge of every tile, sources/map/grid.cpp:141-145, and the Invalid|Building blocking mask). */
int mtd_grid_upload(mtd_ctx *ctx, int32_t W, int32_t H, int32_t ox, int32_t oy, const int16_t *elev, const uint16_t *flags);
int mtd_grid_upload_dev(mtd_ctx *ctx, int32_t W, int32_t H, int32_t ox, int32_t oy, const int16_t *d_elev, const uint16_t *d_flags);
/* placeStructure / clearStructure flag flip (sources/player/buildingsPlacement.cpp:120-130, 143-164):
 * overwrite the flags word of one tile on the device mirror. */
int mtd_grid_set_flags(mtd_ctx *ctx, uint32_t tile, uint16_t flags);
/* Grid::neighborDistance for arbitrary adjacent pairs, computed on the device (parity probe for G4). */
int mtd_neighbor_distances(mtd_ctx *ctx, const uint32_t *a, const uint32_t *b, int32_t n, uint32_t *out);
/* right/down edge weights of every tile as the kernels see them (0 where there is no such neighbour) */
int mtd_edge_weights(mtd_ctx *ctx, uint32_t *right, uint32_t *down);

/* ---- Directions::update (sources/map/paths.cpp:24-78) ----------------------------------------- */
/* nsrc == 1 is the reference; nsrc > 1 seeds every source at distance 0 (synthetic multi-source config).
 * extra_blocked != MTD_M treats that tile as Building without touching the grid -- the trial of
 * placingBuildingBlocksMonsters (sources/player/buildingsPlacement.cpp:18-29). Outputs may be NULL. */
int mtd_directions_update(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t extra_blocked, uint32_t *tiles_out, uint32_t *dist_out);
int mtd_directions_update_dev(mtd_ctx *ctx, const uint32_t *sources /* host */, int32_t nsrc, uint32_t extra_blocked, uint32_t *d_tiles, uint32_t *d_dist);

/* n independent Directions (one target tile each) over the uploaded grid -- what Waypoints::update hands to
 * tasksRunBlocking (sources/map/paths.cpp:82-85, 119) without the path walks. d_tiles (may be NULL) / d_dist = [n][W*H],
 * device. On large grids all layers are relaxed together. This is the unit dealt to the GPUs when layers are sharded. */
int mtd_directions_layers_dev(mtd_ctx *ctx, const uint32_t *targets /* host */, int32_t n, uint32_t *d_tiles, uint32_t *d_dist);

/* ---- Directions::update of ONE large grid split into row stripes, one context (one GPU) per stripe --------
 * (no reference counterpart beyond sources/map/paths.cpp:24-78 itself: the reference is single-threaded per layer;
 * this is the partitioning of SURVEY.md 8e for the 16384^2 configuration.)
 * A stripe owns rows [y0, y1) of a W x H grid. elev_rows / flags_rows hold rows [max(y0-1, 0), min(y1+1, H)), i.e. the
 * owned rows plus one halo row on each inner side (needed for the edge weights across the cut).
 * Protocol, identical on every rank (mazetd_b200/striped.py drives it over torch.distributed):
 *   mtd_stripe_begin_dev(sources = ALL sources as global tile ids, d_dist = [y1-y0][W])   relax locally
 *   repeat { send own first / last row of d_dist to the ranks above / below; receive theirs;
 *            mtd_stripe_exchange_dev(row_above, row_below, &improved) }                  merge halos, relax again
 *   until the sum of `improved` over all ranks is 0 (one all-reduce per sweep)
 *   mtd_stripe_finalize_dev(d_tiles)                                                      next hops, GLOBAL tile ids
 * While the protocol runs, blocked tiles of d_dist read 0 (parked); finalize rewrites them to MTD_M.
 * The distance fixpoint is unique, so the result is bit-identical to one Directions::update over the whole grid. */
int mtd_stripe_upload(mtd_ctx *ctx, int32_t W, int32_t H, int32_t y0, int32_t y1, const int16_t *elev_rows, const uint16_t *flags_rows);
int mtd_stripe_upload_dev(mtd_ctx *ctx, int32_t W, int32_t H, int32_t y0, int32_t y1, const int16_t *d_elev_rows, const uint16_t *d_flags_rows);
int mtd_stripe_begin_dev(mtd_ctx *ctx, const uint32_t *sources /* host */, int32_t nsrc, uint32_t *d_dist);
/* d_row_above = current row y0-1 of the rank above (device, [W]) or NULL at the top; d_row_below likewise row y1.
 * *improved_out = halo cells whose label dropped (0 on every rank = converged). Synchronous. */
int mtd_stripe_exchange_dev(mtd_ctx *ctx, const uint32_t *d_row_above, const uint32_t *d_row_below, uint32_t *improved_out);
int mtd_stripe_finalize_dev(mtd_ctx *ctx, uint32_t *d_tiles);
/* rounds > 0: begin / exchange stop after that many relaxation rounds even if tiles are still queued (0 = run to local
 * convergence). mtd_stripe_pending reports the queued (layer, tile) items; the protocol then ends when the sum of
 * `improved` AND of `pending` over all ranks is 0. Swapping rows more often keeps a stripe from flooding itself with labels
 * that the neighbour stripe is about to undercut. */
int mtd_stripe_round_budget(mtd_ctx *ctx, int32_t rounds);
int mtd_stripe_pending(mtd_ctx *ctx, uint32_t *items_out);

/* Peer-coupled stripes: the same Directions::update (paths.cpp:24-78) as ONE kernel run per rank instead of host-driven
 * sweeps. A border tile stores its boundary row straight into the neighbour rank's halo copy over NVLink and posts the
 * offered key there. Two relaxation kernels (MTD_DEBUG_BIG_ASYNC, same result):
 *   dataflow (default)  no rounds: the neighbour's owner CTA takes the posted key over as a local offer; rank 0 collects
 *                       waves of (idle?, border offers made, border offers consumed) snapshots from all ranks and ends the run
 *                       after two identical quiet waves; every rank publishes its wavefront position for a global priority window
 *   rounds              between rounds every rank stores a phase tag, its queue length and its smallest key into its slot of
 *                       EVERY rank's block and waits on its own block for the others, so all ranks stop in the same round
 *   export         creates this rank's peer block (after mtd_stripe_upload*) and returns its CUDA IPC handle (64 bytes)
 *   connect        maps the blocks of all other ranks from `handles` ([world][64], gathered by the caller; world <= 16)
 *   launch_dev     queues initialisation, the relaxation kernel and the next-hop pass on the context's stream and returns;
 *                  EVERY rank must call it (the others wait 20 s for a rank to enter the run and 4 s at each round barrier,
 *                  then give up -> MTD_ERR_STATE at finish; the barrier counters are then out of step: destroy the contexts)
 *   finish         waits for the stream, reports device status; *rounds_out = rounds the ranks ran */
int mtd_stripe_peer_export(mtd_ctx *ctx, void *handle64);
int mtd_stripe_peer_connect(mtd_ctx *ctx, int32_t rank, int32_t world, const void *handles);
/* the same for stripes that are contexts of THIS process on ONE device (tests: several stripes share a GPU); each kernel is
 * limited to `ctas` CTAs so that all of them are resident together; export every context first */
int mtd_stripe_peer_connect_local(mtd_ctx *ctx, int32_t rank, int32_t world, mtd_ctx *const *all /* [world], all[rank] == ctx */, int32_t ctas);
int mtd_stripe_peer_launch_dev(mtd_ctx *ctx, const uint32_t *sources /* host */, int32_t nsrc, uint32_t *d_dist, uint32_t *d_tiles);
int mtd_stripe_peer_finish(mtd_ctx *ctx, uint32_t *rounds_out);

/* ---- Waypoints::update (sources/map/paths.cpp:82-139) ------------------------------------------ */
/* n Dijkstra layers + n full-path walks + aggregates, one call.
 *   tiles_out, dist_out : [n][W*H]            (may be NULL)
 *   fullpath_buf        : [n][fullpath_cap]   tiles of Waypoint::fullPath, MTD_M padded (may be NULL)
 *   fullpath_len        : [n]  true path lengths (may exceed cap; only cap entries are written)
 *   full_distance       : [n]  Waypoint::fullDistance
 *   agg3                : avgFullDistance, minFullDistance, minDistanceSpawner
 * The layers stay resident on the device for mtd_waypoints_find_dev / monsters / placing_blocks. */
int mtd_waypoints_update(mtd_ctx *ctx, const uint32_t *wp_tiles, int32_t n, uint32_t *tiles_out, uint32_t *dist_out, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3);
int mtd_waypoints_update_dev(mtd_ctx *ctx, const uint32_t *wp_tiles /* host */, int32_t n);
/* Waypoints::find (sources/map/paths.cpp:141-169) evaluated on the device for a batch of queries
 * (host arrays in/out). The adapter keeps the per-call host version on host mirrors, as SURVEY.md 8b asks. */
int mtd_waypoints_find(mtd_ctx *ctx, const uint32_t *pos, const uint32_t *bits, int32_t nq, uint32_t *tile_out, uint32_t *dist_out);
/* device pointers of the resident layers ([n][W*H] each); valid until the next update/upload */
int mtd_waypoints_layers_dev(mtd_ctx *ctx, uint32_t **d_tiles, uint32_t **d_dist);

/* ---- placingBuildingBlocksMonsters (sources/player/buildingsPlacement.cpp:18-42) ---------------- */
/* One trial Dijkstra from waypoint 0 with `candidate` blocked; *blocked_out = 1 if any waypoint tile or any
 * live resident monster (mtd_monsters_upload) becomes unreachable. */
int mtd_placing_blocks(mtd_ctx *ctx, uint32_t candidate, int32_t *blocked_out);

/* ---- procedural generator (sources/map/procedural.cpp:13-70, sources/map/grid.cpp:147-198,
 *      sources/map/paths.cpp:173-201) ------------------------------------------------------------- */
typedef struct mtd_mapgen_params
{
	int32_t W, H;     /* reference: 101, 101 (grid.cpp:150) */
	int32_t ox, oy;   /* reference: 50, 50   (grid.cpp:151) */
	int32_t radius;   /* reference: 45       (grid.cpp:179); < 0 disables the radius test */
	int32_t nwp;      /* reference: 4        (paths.cpp:190) */
	uint32_t fullpath_cap;
} mtd_mapgen_params;

/* Generate maps k = first .. first+nmaps-1 from RNG seed (s1 + k, s2): newProcedural -> newGrid ->
 * newWaypoints (incl. Waypoints::update). All outputs are optional host arrays, map-major:
 *   elev [nmaps][T] int16, flags [nmaps][T] uint16, wp_tiles [nmaps][nwp],
 *   tiles/dist [nmaps][nwp][T], fullpath [nmaps][nwp][cap], fullpath_len/full_distance [nmaps][nwp],
 *   agg [nmaps][3]. */
int mtd_mapgen_batch(mtd_ctx *ctx, const mtd_mapgen_params *p, uint64_t s1, uint64_t s2, int32_t first, int32_t nmaps, int16_t *elev, uint16_t *flags, uint32_t *wp_tiles, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg);
/* same with DEVICE output pointers (all required except tiles/dist/fullpath, which may be NULL); asynchronous */
int mtd_mapgen_batch_dev(mtd_ctx *ctx, const mtd_mapgen_params *p, uint64_t s1, uint64_t s2, int32_t first, int32_t nmaps, int16_t *d_elev, uint16_t *d_flags, uint32_t *d_wp_tiles, uint32_t *d_tiles, uint32_t *d_dist, uint32_t *d_fullpath, uint32_t *d_fullpath_len, uint32_t *d_full_distance, uint32_t *d_agg);

/* Allocation-free variant. NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Waypoints::update + full-path walks of `nmaps` maps that are already on the device: d_elev / d_flags [nmaps][W*H],
 * d_wp_tiles [nmaps][n] in, d_fullpath_len / d_full_distance [nmaps][n] and d_agg [nmaps][3] out (device pointers, all
 * required). The kernel's scratch -- the distance and next-hop layers its walker warps chase, and its map counter -- is a
 * fixed array inside the library image (MTD_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library until exit):
 * the call makes no cudaMalloc / cudaFree and no allocating library call. Limit: 256 + nmaps * n * W * H * 8 bytes must fit the
 * array (e.g. 12 maps of 101 x 101 with 4 waypoints), otherwise MTD_ERR_CAPACITY. Asynchronous on the context's stream like
 * the other _dev calls (kernel status through mtd_synchronize). Every call, from any context, shares the one array per GPU:
 * calls that overlap on a GPU race; the array keeps the last call's layers (its control words are cleared by each call). */
#define MTD_NOALLOC_SCRATCH_BYTES (4u << 20)
int mtd_waypoints_paths_noalloc(mtd_ctx *ctx, int32_t W, int32_t H, int32_t nmaps, int32_t n, const int16_t *d_elev, const uint16_t *d_flags, const uint32_t *d_wp_tiles, uint32_t *d_fullpath_len, uint32_t *d_full_distance, uint32_t *d_agg);
/* newGrid after the elevation sampling (sources/map/grid.cpp:160-196): biome flags, maxSlope / radius invalidation,
 * fillUnreachable (largest 4-connected component, first-discovered wins ties), findBorder, Mana -- on the CALLER's
 * int16 elevations [nmaps][T] (host), flags out [nmaps][T] (host). Uses W, H, ox, oy, radius of *p. This is the part of
 * the generator that the compiled reference pins independently of the noise library. */
int mtd_grid_from_elevations(mtd_ctx *ctx, const mtd_mapgen_params *p, int32_t nmaps, const int16_t *elev, uint16_t *flags_out);
/* raw float32 ProceduralImpl::elevation samples (procedural.cpp:65-70) for seed (s1, s2): parity probe */
int mtd_elevation_samples(mtd_ctx *ctx, uint64_t s1, uint64_t s2, const float *xy, int32_t n, float *out);

/* ---- terrain SDF for marching cubes (the mesh itself stays out of scope) -------------------- */
/* Maker::sdf (sources/map/generate.cpp:85-96) = ProceduralImpl::sdf (sources/map/procedural.cpp:134) + the cliff term of
 * invalid land tiles, sampled at every lattice position of MarchingCubes::updateByPosition
 * (externals/cage/sources/libcore/mesh/marchingCubes.cpp:149-165, 247-257). The game uses box (-55,-15,-55)..(55,15,55) and
 * resolution (220, 60, 220) (generate.cpp:107-117). flags = the generated grid's TileFlags (host, [W*H]); out = host
 * [res[2]][res[1]][res[0]] float32 (x fastest). A headless buffer: what the reference hands to dualmc. */
int mtd_sdf_lattice(mtd_ctx *ctx, uint64_t s1, uint64_t s2, int32_t W, int32_t H, int32_t ox, int32_t oy, const uint16_t *flags, const float *box_a, const float *box_b, const int32_t *res, float *out);

/* ---- monster steering (sources/game/monsters.cpp:55-99, sources/game/game.cpp:15-21) ------------ */
/* SoA mirror of PositionComponent / MovementComponent / MonsterComponent fields the tick touches
 * (sources/game.h:154-186). dotflags: bit0 = affected(Fire), bit1 = affected(Water); NULL = none.
 * damage NULL = 0. alive NULL = all alive. */
int mtd_monsters_upload(mtd_ctx *ctx, int32_t n, const uint32_t *tile, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, const uint32_t *bits, const uint32_t *timeToArrive, const uint8_t *dotflags, const uint32_t *damage, const uint8_t *alive);
int mtd_monsters_download(mtd_ctx *ctx, uint32_t *tile, uint32_t *tileStart, uint32_t *tileEnd, uint32_t *timeStart, uint32_t *timeEnd, uint32_t *bits, uint8_t *alive);
/* one "move monsters" tick at gameTime against the resident waypoint layers. Asynchronous. */
int mtd_monsters_tick_dev(mtd_ctx *ctx, uint32_t gameTime);
/* synchronous variant returning the damage dealt to the player this tick and the number of live monsters */
int mtd_monsters_tick(mtd_ctx *ctx, uint32_t gameTime, int64_t *damage_out, uint32_t *alive_out);
/* MovementComponent::position() for every monster -> [n][3] float32 (device or host output) */
int mtd_monsters_positions_dev(mtd_ctx *ctx, uint32_t gameTime, float *d_out3n);
int mtd_monsters_positions(mtd_ctx *ctx, uint32_t gameTime, float *out3n);
/* SpawningGroup::spawnOne steering state (sources/game/monsterSpawning.cpp:205-216): out3 = tile, bits, timeToArrive */
int mtd_monster_spawn(mtd_ctx *ctx, uint32_t gameTime, float speed, uint32_t *out3);

/* ---- damage-over-time tick: "damage monsters", sources/game/monsters.cpp:101-183 ------------ */
/* Integer state only (the RNG-driven visual effects of :140-149 and the ghost of :176 stay with the caller). The monsters
 * are those of mtd_monsters_upload. duration / damage: [5][n] in DamageTypeEnum order Physical, Fire, Water, Poison, Magic
 * (sources/game.h:18-27, 166-172); resist: [n] DamageTypeFlags bits 0..4 (may be NULL); life: [n] sint64; money: [n]
 * (may be NULL). After this upload the steering tick reads affected(Fire) / affected(Water) from these dots. */
int mtd_monsters_dots_upload(mtd_ctx *ctx, const uint32_t *duration, const uint32_t *damage, const uint8_t *resist, const int64_t *life, const uint32_t *money);
/* one tick: dotsEliminateOposing x2, applyDot x5, life <= 0 -> dead (alive = 0), its money added to *money_out */
int mtd_monsters_dots_tick_dev(mtd_ctx *ctx);
int mtd_monsters_dots_tick(mtd_ctx *ctx, int64_t *money_out, uint32_t *killed_out);
int mtd_monsters_dots_download(mtd_ctx *ctx, uint32_t *duration, uint32_t *damage, int64_t *life, uint8_t *alive);

/* ---- Towers: monster spatial index + attack pass ------------------------------
 * Replaces, per game tick, the "spatial rebuild" listener (sources/game/spatial.cpp:24-34: cage's BVH over
 * MovementComponent::position() of every monster) and the "attacks" listener (sources/game/attacks.cpp:132-160) on the
 * resident monster SoA. The BVH becomes a uniform grid (one bin per map tile, counting sort); every tower is served by one
 * warp: firing delay, applyMods (:40-71), sphere query + class filter (:73-92), mana (:95-105), target picker (:171-216),
 * splash query (:94-98), damage-over-time added to the hit monsters (:120-130).
 * Parity: the SET a query returns and the picker's VALUE order are the reference's; the order in which cage's BVH
 * enumerates a set is not observable, so ties of a value go to the lowest monster index. TargetingEnum::Random depends
 * on that order and on the global RNG stream: it picks candidate number random_words[tower] % count in index order
 * (documented as unpinned). Rendering effects (:107-117) stay on the host.
 * Tower arrays: PositionComponent::tile, DamageComponent {damage, overTime, firingPeriod, firingRange, splashRadius,
 * manaCost, invalidClasses}, AttackComponent {element 0..4, bonus 0..6, targeting 0..6, firingDelay},
 * ManaStorageComponent::mana (sources/game.h:206-245). Needs mtd_monsters_upload + mtd_monsters_dots_upload before. */
int mtd_monsters_class_upload(mtd_ctx *ctx, const uint8_t *monsterClass /* MonsterClassFlags per monster, NULL = none */);
int mtd_towers_upload(mtd_ctx *ctx, int32_t n, const uint32_t *tile, const uint32_t *damage, const uint32_t *overTime, const uint32_t *firingPeriod, const float *firingRange, const float *splashRadius,
					  const uint32_t *manaCost, const uint8_t *invalidClasses, const uint8_t *element, const uint8_t *bonus, const uint8_t *targeting, const uint32_t *firingDelay, const uint32_t *mana);
/* one tick at gameTime: index rebuild + every tower; random_words: host array [n] for Random targeting or NULL. Asynchronous. */
int mtd_towers_attack_dev(mtd_ctx *ctx, uint32_t gameTime, const uint64_t *random_words);
/* tower state after the last tick (any pointer may be NULL): firingDelay, mana, picked monster (-1: did not fire), monsters hit. Synchronises. */
int mtd_towers_download(mtd_ctx *ctx, uint32_t *firingDelay, uint32_t *mana, int32_t *target, uint32_t *hits);

/* ---- C3 over several GPUs (monsters sharded by index, the 326 KB of layers broadcast per accepted event) ----
 * After another rank's layers were copied INTO this context's resident layer buffers (the device pointers of
 * mtd_waypoints_layers_dev, e.g. by ncclBroadcast), rebuild what Waypoints::find reads beside them (the inter-waypoint table). */
int mtd_waypoints_layers_refresh(mtd_ctx *ctx);

#ifdef __cplusplus
}
#endif

#endif
