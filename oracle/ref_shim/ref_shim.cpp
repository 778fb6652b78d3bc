/*
 * ref_shim.cpp -- TEST INFRASTRUCTURE (oracle/_ref build only; never linked into the product).
 *
 * Lets the reference's OWN translation units
 *     sources/map/paths.cpp, sources/map/grid.cpp, sources/map/procedural.cpp,
 *     externals/cage/sources/libcore/math/random.cpp, externals/cage/sources/libcore/noiseFunction.cpp
 * be compiled unmodified, in place under /root/reference, into oracle/_ref/libmazetd_ref.so
 * (recipe: oracle/Makefile). This file supplies
 *   (1) the handful of cage-core symbols those objects leave undefined (memory arena, Holder
 *       refcount, sqrt/sin/cos, task pool, logging/exception plumbing) -- own minimal code, and
 *   (2) a plain C API (ref_*) that builds Grid / Directions / Waypoints objects from caller arrays
 *       and runs the reference functions on them, so python tests can compare bit for bit.
 * The monster tick (sources/game/monsters.cpp:60-98) and MovementComponent::position()
 * (sources/game/game.cpp:15-21) cannot be linked (they need the ECS and the engine); they are
 * restated here over SoA arrays but call the COMPILED reference Waypoints::find,
 * Grid::neighborDistance, Grid::center and cage's own inline saturate/interpolate.
 */
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cage-core/math.h>
#include <cage-core/memoryArena.h>
#include <cage-core/pointerRangeHolder.h>
#include <cage-core/random.h>
#include <cage-core/tasks.h>

#include "map/generate.h"

// ---------------------------------------------------------------------------------------------
// (1) cage-core symbols
// ---------------------------------------------------------------------------------------------
namespace cage
{
	namespace
	{
		struct ShimArena
		{
			void *allocate(uintPtr size, uintPtr alignment)
			{
				if (alignment < sizeof(void *))
					alignment = sizeof(void *);
				void *p = nullptr;
				if (posix_memalign(&p, alignment, size ? size : 1) != 0)
				{
					std::fprintf(stderr, "ref_shim: out of memory\n");
					std::abort();
				}
				return p;
			}
			void deallocate(void *ptr) { std::free(ptr); }
			void flush() { std::abort(); }
			MemoryArena arena = MemoryArena(this);
		};
	}

	void *MemoryArena::allocate(uintPtr size, uintPtr alignment)
	{
		return stub->alloc(inst, size, alignment);
	}

	void MemoryArena::deallocate(void *ptr)
	{
		stub->dealloc(inst, ptr);
	}

	MemoryArena &systemMemory()
	{
		static ShimArena *a = new ShimArena();
		return a->arena;
	}

	namespace privat
	{
		void HolderControlBase::inc()
		{
			__atomic_add_fetch(&counter, (uint32)1, __ATOMIC_RELAXED);
		}

		void HolderControlBase::dec()
		{
			if (__atomic_sub_fetch(&counter, (uint32)1, __ATOMIC_SEQ_CST) == 0)
				deleter(deletee);
		}

		uint32 toStringImpl(char *s, uint32 n, const char *src)
		{
			const size_t l = std::strlen(src);
			if (l > n)
				std::abort();
			std::memcpy(s, src, l + 1);
			return (uint32)l;
		}

		uint32 toString(char *s, uint32 n, uint32 v)
		{
			return (uint32)std::snprintf(s, n, "%u", v);
		}

		uint32 toString(char *s, uint32 n, uint64 v)
		{
			return (uint32)std::snprintf(s, n, "%llu", (unsigned long long)v);
		}

		uint64 makeLog(const std::source_location &, SeverityEnum, StringPointer, const String &, bool, bool) noexcept
		{
			return 0; // the oracle is silent
		}

		void generateRandomData(uint8 *target, uint32 size)
		{
			// default-constructed generators are always re-seeded explicitly by ref_rng_seed()
			for (uint32 i = 0; i < size; i++)
				target[i] = (uint8)(0x9e + 37 * i);
		}
	}

	namespace detail
	{
		void *memset(void *destination, int value, uintPtr num) { return std::memset(destination, value, num); }
		void *memcpy(void *destination, const void *source, uintPtr num) { return std::memcpy(destination, source, num); }
		void *memmove(void *destination, const void *source, uintPtr num) { return std::memmove(destination, source, num); }
		int memcmp(const void *ptr1, const void *ptr2, uintPtr num) { return std::memcmp(ptr1, ptr2, num); }

		void logCurrentCaughtException() noexcept {}

		void irrecoverableError(StringPointer explanation) noexcept
		{
			std::fprintf(stderr, "ref_shim: irrecoverable: %s\n", (const char *)explanation);
			std::abort();
		}
	}

	Exception::Exception(const std::source_location &location, SeverityEnum severity, StringPointer message) noexcept : location(location), message(message), severity(severity) {}
	Exception::~Exception() {}
	void Exception::makeLog() const noexcept {}
	void Exception::log() const {}

	Real sqrt(Real x)
	{
		return std::sqrt(x.value); // externals/cage/sources/libcore/math/scalars.cpp:99-102
	}

	Real sin(Rads x)
	{
		return std::sin(Real(x).value);
	}

	Real cos(Rads x)
	{
		return std::cos(Real(x).value);
	}

	Quat::Quat(Vec3, Rads)
	{
		std::abort(); // randomDirectionQuat is not on the path
	}

	Vec3 colorHsvToRgb(const Vec3 &)
	{
		std::abort(); // texturing only (sources/map/procedural.cpp:79-132), not on the path
	}

	// One thread per invocation, like the reference pool does when invocations <= cores
	// (externals/cage/sources/libcore/concurrent/tasks.cpp:173-186, 275-278).
	static std::atomic<int> shimTaskThreads{ 1 };

	void tasksRunBlocking(StringPointer, Delegate<void(uint32)> function, uint32 invocations)
	{
		if (shimTaskThreads.load() <= 1 || invocations <= 1)
		{
			for (uint32 i = 0; i < invocations; i++)
				function(i);
			return;
		}
		std::vector<std::thread> thrs;
		thrs.reserve(invocations);
		for (uint32 i = 0; i < invocations; i++)
			thrs.emplace_back([function, i]() { function(i); });
		for (auto &t : thrs)
			t.join();
	}
}

// ---------------------------------------------------------------------------------------------
// (2) C API over the reference's objects
// ---------------------------------------------------------------------------------------------
using namespace cage;
using namespace mazetd;

namespace
{
	Holder<Grid> makeGrid(int W, int H, int ox, int oy, const sint16 *elev, const uint16 *flags)
	{
		Holder<Grid> g = systemMemory().createHolder<Grid>();
		g->resolution = Vec2i(W, H);
		g->gridOffset = Vec2i(ox, oy);
		const uint32 total = (uint32)W * (uint32)H;
		{
			PointerRangeHolder<const sint16> vec;
			vec.resize(total);
			if (elev)
				std::memcpy((void *)vec.data(), elev, total * sizeof(sint16));
			g->elevations = vec;
		}
		{
			PointerRangeHolder<TileFlags> vec;
			vec.resize(total);
			if (flags)
				std::memcpy((void *)vec.data(), flags, total * sizeof(uint16));
			g->flags = vec;
		}
		return g;
	}

	struct RefMap
	{
		Holder<Grid> grid;
		Holder<Waypoints> wps;
	};
}

extern "C"
{
	void ref_set_task_threads(int n)
	{
		cage::shimTaskThreads.store(n);
	}

	// ---- RNG (externals/cage/sources/libcore/math/random.cpp) ----
	void ref_rng_seed(uint64_t s1, uint64_t s2)
	{
		detail::randomGenerator() = RandomGenerator(s1, s2);
	}

	void ref_rng_state(uint64_t *out2)
	{
		out2[0] = detail::randomGenerator().s[0];
		out2[1] = detail::randomGenerator().s[1];
	}

	uint64_t ref_rng_next()
	{
		return detail::randomGenerator().next();
	}

	float ref_rng_chance()
	{
		return detail::randomGenerator().randomChance().value;
	}

	uint32_t ref_rng_range_u32(uint32_t a, uint32_t b)
	{
		return randomRange(a, b);
	}

	// known-answer cases of externals/cage/sources/test-core/random.cpp:47-78
	void ref_rng_kat_range4i(uint64_t s1, uint64_t s2, int32_t a, int32_t b, int32_t *out4)
	{
		RandomGenerator g(s1, s2);
		const Vec4i v = g.randomRange4i(a, b);
		for (int i = 0; i < 4; i++)
			out4[i] = v[i];
	}

	void ref_rng_kat_range4(uint64_t s1, uint64_t s2, float a, float b, float *out4)
	{
		RandomGenerator g(s1, s2);
		const Vec4 v = g.randomRange4(a, b);
		for (int i = 0; i < 4; i++)
			out4[i] = v[i].value;
	}

	// ---- fixed point helpers (sources/common.h:15-29) ----
	int32_t ref_rtos32(float v)
	{
		return rtos32(v);
	}

	int16_t ref_rtos16(float v)
	{
		return rtos16(v);
	}

	float ref_stor(int32_t v)
	{
		return stor(v).value;
	}

	// ---- Grid geometry (sources/map/grid.cpp:100-145) ----
	uint32_t ref_grid_index_i(int W, int H, int ox, int oy, int x, int y)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, nullptr, nullptr);
		return g->index(Vec2i(x, y));
	}

	uint32_t ref_grid_index_f(int W, int H, int ox, int oy, float x, float y)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, nullptr, nullptr);
		return g->index(Vec2(x, y));
	}

	void ref_grid_position(int W, int H, int ox, int oy, uint32_t idx, int32_t *out2)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, nullptr, nullptr);
		const Vec2i p = g->position(idx);
		out2[0] = p[0];
		out2[1] = p[1];
	}

	void ref_grid_centers(int W, int H, int ox, int oy, const int16_t *elev, const uint32_t *idx, int n, float *out3n)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, elev, nullptr);
		for (int i = 0; i < n; i++)
		{
			const Vec3 c = g->center(idx[i]);
			for (int k = 0; k < 3; k++)
				out3n[i * 3 + k] = c[k].value;
		}
	}

	void ref_neighbor_distances(int W, int H, int ox, int oy, const int16_t *elev, const uint32_t *a, const uint32_t *b, int n, uint32_t *out)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, elev, nullptr);
		for (int i = 0; i < n; i++)
			out[i] = g->neighborDistance(a[i], b[i]);
	}

	// right-edge and down-edge weights for every tile (0 where the neighbour is outside the grid)
	void ref_edge_weights(int W, int H, const int16_t *elev, uint32_t *right, uint32_t *down)
	{
		Holder<Grid> g = makeGrid(W, H, 0, 0, elev, nullptr);
		for (int y = 0; y < H; y++)
			for (int x = 0; x < W; x++)
			{
				const uint32 i = (uint32)y * W + x;
				right[i] = x + 1 < W ? g->neighborDistance(i, i + 1) : 0;
				down[i] = y + 1 < H ? g->neighborDistance(i, i + W) : 0;
			}
	}

	// ---- Directions::update (sources/map/paths.cpp:24-78) ----
	void ref_directions_update(int W, int H, int ox, int oy, const int16_t *elev, const uint16_t *flags, uint32_t tile, uint32_t *tiles_out, uint32_t *dist_out)
	{
		Holder<Grid> g = makeGrid(W, H, ox, oy, elev, flags);
		Holder<Directions> d = systemMemory().createHolder<Directions>();
		d->grid = g.share();
		d->tile = tile;
		d->update();
		const size_t total = (size_t)W * H;
		std::memcpy(tiles_out, d->tiles.data(), total * 4);
		std::memcpy(dist_out, d->distances.data(), total * 4);
	}

	// ---- Waypoints (sources/map/paths.cpp:82-169) ----
	void *ref_map_create(int W, int H, int ox, int oy, const int16_t *elev, const uint16_t *flags, const uint32_t *wp_tiles, int n)
	{
		RefMap *mp = new RefMap();
		mp->grid = makeGrid(W, H, ox, oy, elev, flags);
		mp->wps = systemMemory().createHolder<Waypoints>();
		PointerRangeHolder<Holder<Waypoints::Waypoint>> vec;
		vec.reserve(n);
		for (int i = 0; i < n; i++)
		{
			Holder<Waypoints::Waypoint> p = systemMemory().createHolder<Waypoints::Waypoint>();
			p->grid = mp->grid.share();
			p->tile = wp_tiles[i];
			vec.push_back(std::move(p));
		}
		mp->wps->waypoints = vec;
		return mp;
	}

	void ref_map_destroy(void *h)
	{
		delete (RefMap *)h;
	}

	void ref_map_set_flags(void *h, const uint16_t *flags)
	{
		RefMap *mp = (RefMap *)h;
		std::memcpy((void *)mp->grid->flags.data(), flags, mp->grid->flags.size() * 2);
	}

	void ref_map_get_grid(void *h, int32_t *whoo, int16_t *elev, uint16_t *flags)
	{
		RefMap *mp = (RefMap *)h;
		if (whoo)
		{
			whoo[0] = mp->grid->resolution[0];
			whoo[1] = mp->grid->resolution[1];
			whoo[2] = mp->grid->gridOffset[0];
			whoo[3] = mp->grid->gridOffset[1];
		}
		if (elev)
			std::memcpy(elev, mp->grid->elevations.data(), mp->grid->elevations.size() * 2);
		if (flags)
			std::memcpy(flags, mp->grid->flags.data(), mp->grid->flags.size() * 2);
	}

	int ref_map_waypoint_count(void *h)
	{
		return (int)((RefMap *)h)->wps->waypoints.size();
	}

	void ref_map_waypoint_tiles(void *h, uint32_t *out)
	{
		RefMap *mp = (RefMap *)h;
		for (uint32 i = 0; i < mp->wps->waypoints.size(); i++)
			out[i] = mp->wps->waypoints[i]->tile;
	}

	void ref_waypoints_update(void *h)
	{
		((RefMap *)h)->wps->update();
	}

	void ref_waypoints_get_layer(void *h, int i, uint32_t *tiles_out, uint32_t *dist_out)
	{
		RefMap *mp = (RefMap *)h;
		const auto &w = mp->wps->waypoints[i];
		if (tiles_out)
			std::memcpy(tiles_out, w->tiles.data(), w->tiles.size() * 4);
		if (dist_out)
			std::memcpy(dist_out, w->distances.data(), w->distances.size() * 4);
	}

	uint32_t ref_waypoints_fullpath(void *h, int i, uint32_t *buf, uint32_t cap, uint32_t *full_distance)
	{
		RefMap *mp = (RefMap *)h;
		const auto &w = mp->wps->waypoints[i];
		const uint32 n = (uint32)w->fullPath.size();
		if (buf)
			std::memcpy(buf, w->fullPath.data(), (size_t)(n < cap ? n : cap) * 4);
		if (full_distance)
			*full_distance = w->fullDistance;
		return n;
	}

	void ref_waypoints_aggregates(void *h, uint32_t *out3)
	{
		RefMap *mp = (RefMap *)h;
		out3[0] = mp->wps->avgFullDistance;
		out3[1] = mp->wps->minFullDistance;
		out3[2] = mp->wps->minDistanceSpawner;
	}

	void ref_waypoints_find(void *h, uint32_t pos, uint32_t bits, uint32_t *out2)
	{
		const auto r = ((RefMap *)h)->wps->find(pos, bits);
		out2[0] = r.tile;
		out2[1] = r.distance;
	}

	// ---- map generation (sources/map/generate.cpp:155-165 minus meshes) ----
	// Seeds the calling thread's generator, then newProcedural -> newGrid -> newWaypoints.
	void *ref_mapgen(uint64_t s1, uint64_t s2)
	{
		detail::randomGenerator() = RandomGenerator(s1, s2);
		RefMap *mp = new RefMap();
		Holder<Procedural> procedural = newProcedural();
		mp->grid = newGrid(procedural.share());
		mp->wps = newWaypoints(mp->grid.share());
		return mp;
	}

	// Batch of independent maps k = first .. first+nmaps-1 with seed (s1 + k, s2), `threads` host threads, one map
	// per thread at a time (the reference generates one map per pool thread, sources/map/generate.cpp:193).
	// Each map runs the reference's newProcedural -> newGrid -> newWaypoints (incl. Waypoints::update) and is
	// destroyed again; returns a checksum of the aggregates so the work cannot be optimised away.
	uint64_t ref_mapgen_batch(uint64_t s1, uint64_t s2, int first, int nmaps, int threads)
	{
		if (threads < 1)
			threads = 1;
		std::atomic<int> next{ 0 };
		std::atomic<uint64_t> checksum{ 0 };
		auto worker = [&]()
		{
			uint64_t cs = 0;
			for (;;)
			{
				const int k = next.fetch_add(1);
				if (k >= nmaps)
					break;
				RefMap *mp = (RefMap *)ref_mapgen(s1 + (uint64_t)(first + k), s2);
				cs += (uint64_t)mp->wps->avgFullDistance + ((uint64_t)mp->wps->minFullDistance << 20) + ((uint64_t)mp->wps->minDistanceSpawner << 40);
				delete mp;
			}
			checksum.fetch_add(cs);
		};
		std::vector<std::thread> thrs;
		for (int t = 1; t < threads; t++)
			thrs.emplace_back(worker);
		worker();
		for (auto &t : thrs)
			t.join();
		return checksum.load();
	}

	// newGrid only, on caller-provided RNG seed; no waypoints
	void *ref_mapgen_grid_only(uint64_t s1, uint64_t s2)
	{
		detail::randomGenerator() = RandomGenerator(s1, s2);
		RefMap *mp = new RefMap();
		Holder<Procedural> procedural = newProcedural();
		mp->grid = newGrid(procedural.share());
		return mp;
	}

	// raw float elevation samples of ProceduralImpl::elevation (sources/map/procedural.cpp:65-70)
	void ref_elevation_samples(uint64_t s1, uint64_t s2, const float *xy, int n, float *out)
	{
		detail::randomGenerator() = RandomGenerator(s1, s2);
		Holder<Procedural> procedural = newProcedural();
		for (int i = 0; i < n; i++)
			out[i] = procedural->elevation(Vec2(xy[i * 2 + 0], xy[i * 2 + 1])).value;
	}

	// ---- terrain SDF over the marching-cubes lattice ----
	// Maker::sdf (sources/map/generate.cpp:85-96) and MarchingCubesCreateConfig::position
	// (externals/cage/sources/libcore/mesh/marchingCubes.cpp:247-257) cannot be linked (mesh library, dualmc); they are
	// restated here with cage's own Vec3 arithmetic and call the COMPILED reference Procedural::sdf, newGrid and Grid::index.
	// Storage order as MarchingCubes::updateByPosition (marchingCubes.cpp:149-165). flags_out: the grid's flags after newGrid.
	void ref_sdf_lattice(uint64_t s1, uint64_t s2, const float *a, const float *b, const int32_t *res, float *out, uint16_t *flags_out)
	{
		detail::randomGenerator() = RandomGenerator(s1, s2);
		Holder<Procedural> procedural = newProcedural();
		Holder<Grid> grid = newGrid(procedural.share());
		if (flags_out)
			std::memcpy(flags_out, grid->flags.data(), grid->flags.size() * sizeof(uint16));
		const Vec3 boxA(a[0], a[1], a[2]), boxB(b[0], b[1], b[2]);
		const Vec3i resolution(res[0], res[1], res[2]);
		size_t o = 0;
		for (uint32 z = 0; z < (uint32)res[2]; z++)
			for (uint32 y = 0; y < (uint32)res[1]; y++)
				for (uint32 x = 0; x < (uint32)res[0]; x++)
				{
					const Vec3 f = (Vec3(x, y, z) - 2) / (Vec3(resolution) - 5);
					const Vec3 pos = (boxB - boxA) * f + boxA;
					Real d = procedural->sdf(pos);
					const uint32 i = grid->index(pos);
					if (i != m)
					{
						const TileFlags fl = grid->flags[i];
						if (any(fl & TileFlags::Invalid) && none(fl & TileFlags::Water))
							d += saturate(find(pos[1], -7.5, -6.5));
					}
					out[o++] = d.value;
				}
	}

	// ---- monster steering: restatement over SoA calling the compiled reference map functions ----
	// follows sources/game/monsters.cpp:60-98 statement by statement.
	// alive[i]==0 entries are skipped; a monster that completes all waypoints gets alive=0 and its damage is
	// added to *damage_sum (playerHealth -= mo.damage, monsters.cpp:73).
	void ref_monsters_tick(void *h, uint32_t gameTime, int n, uint32_t *tile, uint32_t *tileStart, uint32_t *tileEnd, uint32_t *timeStart, uint32_t *timeEnd, uint32_t *bits, const uint32_t *timeToArrive, const uint8_t *dotflags /*bit0 fire, bit1 water*/, const uint32_t *damage, uint8_t *alive, int64_t *damage_sum)
	{
		RefMap *mp = (RefMap *)h;
		const uint32 nw = (uint32)mp->wps->waypoints.size();
		for (int i = 0; i < n; i++)
		{
			if (!alive[i])
				continue;
			if (gameTime < timeEnd[i])
				continue;
			for (uint32 w = 0; w < nw; w++)
				if (mp->wps->waypoints[w]->tile == tile[i])
					bits[i] |= 1u << w;
			if ((uint32)__builtin_popcount(bits[i]) == nw)
			{
				*damage_sum += damage ? damage[i] : 0;
				alive[i] = 0;
				continue;
			}
			const auto go = mp->wps->find(tile[i], bits[i]);
			tileStart[i] = tile[i];
			tileEnd[i] = tile[i] = go.tile;
			timeStart[i] = gameTime;
			if (timeToArrive[i] > gameTime)
			{
				const uint64 distNext = mp->grid->neighborDistance(tileStart[i], tileEnd[i]);
				timeEnd[i] = numeric_cast<uint32>(gameTime + (timeToArrive[i] - gameTime) * distNext / go.distance);
			}
			timeEnd[i] = max(timeEnd[i], gameTime + 1);
			const bool fire = dotflags && (dotflags[i] & 1), water = dotflags && (dotflags[i] & 2);
			if (water && !fire)
				timeEnd[i] = gameTime + (timeEnd[i] - gameTime) * 2;
			if (fire && !water)
				timeEnd[i] = gameTime + (timeEnd[i] - gameTime) / 2;
			timeEnd[i] = max(timeEnd[i], gameTime + 1);
		}
	}

	// sources/game/monsterSpawning.cpp:205-216
	void ref_monster_spawn(void *h, uint32_t gameTime, float speed, uint32_t *out /*tile,bits,timeToArrive*/)
	{
		RefMap *mp = (RefMap *)h;
		const uint32 position = mp->wps->waypoints[mp->wps->minDistanceSpawner]->tile;
		const uint32 bits = 1u << mp->wps->minDistanceSpawner;
		const Real sp = speed;
		out[0] = position;
		out[1] = bits;
		out[2] = gameTime + numeric_cast<uint32>(30 * stor(mp->wps->find(position, bits).distance) / sp);
	}

	// sources/game/game.cpp:15-21
	void ref_monsters_position(void *h, uint32_t gameTime, int n, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, float *out3n)
	{
		RefMap *mp = (RefMap *)h;
		for (int i = 0; i < n; i++)
		{
			const Vec3 ca = mp->grid->center(tileStart[i]);
			const Vec3 cb = mp->grid->center(tileEnd[i]);
			const Real fac = saturate(Real(gameTime - (sint64)timeStart[i]) / Real(timeEnd[i] - (sint64)timeStart[i]));
			const Vec3 p = interpolate(ca, cb, fac);
			for (int k = 0; k < 3; k++)
				out3n[i * 3 + k] = p[k].value;
		}
	}
}
