/*
 * mazetd_oracle.h -- TEST INFRASTRUCTURE. CPU restatement (plain C) of mazetd's flow-field hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker / reported CPU baseline. The product (mazetd_b200/) never links it.
 *
 * Pinned against:
 *   - the reference's own golden vectors: GridTester (sources/map/grid.cpp:214-251), PRNG known answers
 *     (externals/cage/sources/test-core/random.cpp:47-78), numeric_cast truncation
 *     (externals/cage/sources/test-core/numericTypes.cpp:49-60);
 *   - outputs of the reference itself compiled here (oracle/_ref/libmazetd_ref.so, see oracle/Makefile):
 *     tests/test_oracle_vs_ref.py compares every function below bit for bit on seeded inputs, and
 *     tests/golden/ holds committed vectors generated from it (tests/golden/make_golden.py).
 * Noise (FastNoiseLite, third party, source absent): PARITY UNPINNED, see oracle/fnl_subset.h.
 */
#ifndef MTD_ORACLE_H
#define MTD_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MO_M 0xFFFFFFFFu /* cage `m` for uint32 (externals/cage/sources/include/cage-core/core.h:237-259) */

/* TileFlags (sources/grid.h:15-28) */
enum
{
	MO_FLAG_INVALID = 1u << 0,
	MO_FLAG_BORDER = 1u << 1,
	MO_FLAG_WAYPOINT = 1u << 2,
	MO_FLAG_WATER = 1u << 5,
	MO_FLAG_SUN = 1u << 6,
	MO_FLAG_WIND = 1u << 7,
	MO_FLAG_SNOW = 1u << 8,
	MO_FLAG_MANA = 1u << 13,
	MO_FLAG_BUILDING = 1u << 14,
	MO_FLAG_TRAP = 1u << 15,
};

/* ---- RNG: xoroshiro128** (externals/cage/sources/libcore/math/random.cpp:27-98) ---- */
typedef struct mo_rng
{
	uint64_t s[2];
} mo_rng;
uint64_t mo_rng_next(mo_rng *g);
double mo_rng_chance_double(mo_rng *g);
float mo_rng_chance(mo_rng *g);
uint32_t mo_rng_range_u32(mo_rng *g, uint32_t min, uint32_t max);
int32_t mo_rng_range_i32(mo_rng *g, int32_t min, int32_t max);
float mo_rng_range_f(mo_rng *g, float min, float max);

/* ---- fixed point (sources/common.h:15-29) ---- */
int16_t mo_rtos16(float v);
int32_t mo_rtos32(float v);
float mo_stor(int32_t v);

/* ---- Grid (sources/grid.h:38-53, sources/map/grid.cpp:100-145) ---- */
typedef struct mo_grid
{
	int32_t W, H;   /* resolution */
	int32_t ox, oy; /* gridOffset */
	const int16_t *elev;
	const uint16_t *flags;
} mo_grid;
uint32_t mo_grid_index_i(const mo_grid *g, int32_t x, int32_t y);
uint32_t mo_grid_index_f(const mo_grid *g, float x, float y);
void mo_grid_position(const mo_grid *g, uint32_t idx, int32_t *out2);
void mo_grid_center(const mo_grid *g, uint32_t idx, float *out3);
uint32_t mo_neighbor_distance(const mo_grid *g, uint32_t a, uint32_t b);
void mo_edge_weights(const mo_grid *g, uint32_t *right, uint32_t *down);

/* ---- Directions::update (sources/map/paths.cpp:24-78); nsrc > 1 is the multi-source extension
 *      (all sources start at distance 0), used only by the synthetic C4 config ---- */
int mo_directions_update(const mo_grid *g, const uint32_t *sources, int nsrc, uint32_t *tiles, uint32_t *dist);

/* ---- Waypoints (sources/map/paths.cpp:82-169) ---- */
typedef struct mo_waypoints
{
	const mo_grid *grid;
	int n;
	const uint32_t *wp_tiles; /* [n] */
	const uint32_t *tiles;    /* [n][T] */
	const uint32_t *dist;     /* [n][T] */
} mo_waypoints;
void mo_waypoints_find(const mo_waypoints *w, uint32_t pos, uint32_t bits, uint32_t *out2 /* tile, distance */);
/* returns path length; writes min(len, cap) tiles */
uint32_t mo_waypoints_fullpath(const mo_waypoints *w, int spawner, uint32_t *buf, uint32_t cap, uint32_t *full_distance);
/* out3 = avgFullDistance, minFullDistance, minDistanceSpawner (paths.cpp:122-138) */
void mo_waypoints_aggregates(const uint32_t *full_distance, int n, uint32_t *out3);
/* whole Waypoints::update: n Dijkstras (optionally on `threads` pthreads like tasksRunBlocking), n walks, aggregates */
int mo_waypoints_update(const mo_grid *g, const uint32_t *wp_tiles, int n, int threads, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3);

/* ---- procedural map generator (sources/map/procedural.cpp:13-70, sources/map/grid.cpp:12-198,
 *      sources/map/paths.cpp:173-201) ---- */
typedef struct mo_procedural
{
	float off_x, off_y;    /* elevationOffset */
	uint32_t seeds[5];     /* elevation, water, grass, dirt, roughness noise seeds in draw order */
	float fractal_bounding;
} mo_procedural;
void mo_procedural_init(mo_procedural *p, mo_rng *g);
float mo_elevation(const mo_procedural *p, float x, float y);
/* Maker::sdf at every marching-cubes lattice position (generate.cpp:85-96, marchingCubes.cpp:149-165, 247-257) */
void mo_sdf_lattice(const mo_procedural *p, const mo_grid *g, const float *a, const float *b, const int32_t *res, float *out);
/* newGrid for an arbitrary resolution/offset/radius (reference: 101, 50, 45) */
void mo_new_grid(const mo_procedural *p, int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, int16_t *elev, uint16_t *flags, uint32_t *valid_tiles);
/* the integer stages of newGrid on caller-provided elevations (flags thresholds + slope + radius) */
void mo_grid_flags_from_elev(int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, const int16_t *elev, uint16_t *flags);
void mo_fill_unreachable(int32_t W, int32_t H, uint16_t *flags);
void mo_find_border(int32_t W, int32_t H, uint16_t *flags);
uint32_t mo_add_mana(int32_t W, int32_t H, uint16_t *flags);
/* newWaypoints' spawn selection: n draws with rejection; sets the Waypoint flag */
void mo_pick_waypoints(mo_rng *g, int32_t W, int32_t H, uint16_t *flags, int n, uint32_t *wp_tiles);
/* full pipeline for one seed: newProcedural -> newGrid -> newWaypoints(+update). All outputs caller-owned. */
int mo_mapgen(uint64_t s1, uint64_t s2, int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, int nwp, int16_t *elev, uint16_t *flags, uint32_t *wp_tiles, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3);
/* CPU baseline helper: nmaps seeds (s1+k, s2) on `threads` pthreads, results discarded except a checksum */
uint64_t mo_mapgen_batch_checksum(uint64_t s1, uint64_t s2, int nmaps, int threads);

/* ---- monster steering (sources/game/monsters.cpp:60-98, sources/game/monsterSpawning.cpp:205-216,
 *      sources/game/game.cpp:15-21) over SoA arrays ---- */
typedef struct mo_monsters
{
	int n;
	uint32_t *tile;
	uint32_t *tileStart, *tileEnd, *timeStart, *timeEnd;
	uint32_t *bits;
	const uint32_t *timeToArrive;
	const uint8_t *dotflags; /* bit0 = affected(Fire), bit1 = affected(Water); may be NULL */
	const uint32_t *damage;  /* may be NULL */
	uint8_t *alive;
} mo_monsters;
void mo_monsters_tick(const mo_waypoints *w, uint32_t gameTime, mo_monsters *ms, int64_t *damage_sum);
/* "damage monsters" (sources/game/monsters.cpp:101-183), integer state only; dot arrays [5][n], Physical..Magic */
void mo_monsters_dots_tick(int n, uint32_t *dot_duration, uint32_t *dot_damage, const uint8_t *resist, int64_t *life, const uint32_t *money, uint8_t *alive, int64_t *money_sum, uint32_t *killed);
void mo_monster_spawn(const mo_waypoints *w, uint32_t minDistanceSpawner, uint32_t gameTime, float speed, uint32_t *out3 /* tile, bits, timeToArrive */);
void mo_monsters_position(const mo_grid *g, uint32_t gameTime, int n, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, float *out3n);
/* trial of sources/player/buildingsPlacement.cpp:18-42: 1 if the candidate building would disconnect a waypoint or a monster */
int mo_placing_blocks(const mo_grid *g, const uint32_t *wp_tiles, int n, uint32_t candidate, const uint32_t *monster_tiles, const uint8_t *monster_alive, int nmonsters, uint32_t *scratch_tiles, uint32_t *scratch_dist);

/* "attacks" (sources/game/attacks.cpp:132-160) with the monster queries of sources/game/spatial.cpp:24-34 done by brute
 * force: the SET a sphere query returns is what is restated (externals/cage/sources/libcore/geometry/geometry.cpp:82-86,
 * 471-474: sqrt(dx*dx + dz*dz) <= radius on the y-flattened positions); the ORDER cage's BVH enumerates it in is not
 * observable, so ties of a picker's value go to the LOWEST monster index here. TargetingEnum::Random (attacks.cpp:216-228)
 * depends on that order and on the global RNG: picks candidate number `random_words[tower] % count` in index order (unpinned).
 * Arrays: towers SoA (DamageComponent / AttackComponent / PositionComponent / ManaStorageComponent fields, sources/game.h:206-245),
 * monsters SoA; dot arrays [5][nm]. target_out: picked monster per tower or -1; hits_out: monsters that received the dot. */
typedef struct
{
	int n;
	const uint32_t *tile;
	const uint32_t *damage, *overTime, *firingPeriod, *manaCost;
	const float *firingRange, *splashRadius;
	const uint8_t *invalidClasses, *element, *bonus, *targeting;
	uint32_t *firingDelay, *mana;
	const uint64_t *random_words; /* may be NULL (0) */
	int32_t *target_out;
	uint32_t *hits_out;
} mo_towers;
void mo_towers_attack(const mo_grid *g, uint32_t gameTime, mo_towers *tw, int nm, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, const uint8_t *alive,
					  const uint8_t *monsterClass, const uint32_t *timeToArrive, const int64_t *life, uint32_t *dot_duration, uint32_t *dot_damage);

#ifdef __cplusplus
}
#endif

#endif
