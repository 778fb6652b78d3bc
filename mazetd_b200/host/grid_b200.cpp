// grid_b200.cpp -- link-time replacement for mazetd's sources/map/grid.cpp behind the unchanged sources/grid.h.
//
//   Grid::index / position / center / up / neighborDistance   reference sources/map/grid.cpp:100-145
//       tiny per-call geometry used all over the game: they stay on the host and read the host arrays.
//       The float32 operation order of center() and neighborDistance() is kept, because distances are compared bit for bit.
//   newGrid(Holder<Procedural>)                                 reference sources/map/grid.cpp:147-198
//       ONE device call when the Procedural came from the replacement newProcedural() (procedural_b200.cpp): it carries the
//       generator state its constructor started from, the device repeats the seven draws and samples the elevations
//       itself (mtd_mapgen_batch with no waypoints: noise -> int16, biome flags, slope / radius invalidation, largest
//       connected component, border, mana; grid.cpp:153-196). A foreign Procedural is an opaque virtual interface
//       (sources/map/generate.h:25-31): its elevations are sampled through it on the CPU and only the stages after
//       them run on the GPU (mtd_grid_from_elevations).
#include "map/generate.h"
#include "mtd_host.h"
#include "procedural_b200.h"

#include <cage-core/pointerRangeHolder.h>

namespace mazetd
{
	namespace
	{
		constexpr sint32 MapSize = 101; // tiles per side of the generated map
		constexpr sint32 MapOrigin = 50; // grid coordinates of world (0, 0)
		constexpr sint32 MapRadius = 45; // tiles further from the origin are outside the playable island
	}

	uint32 Grid::index(Vec2i pos) const
	{
		const sint32 x = pos[0] + gridOffset[0];
		const sint32 y = pos[1] + gridOffset[1];
		const bool inside = x >= 0 && y >= 0 && x < resolution[0] && y < resolution[1];
		return inside ? uint32(y) * uint32(resolution[0]) + uint32(x) : m;
	}

	uint32 Grid::index(Vec2 pos) const
	{
		// round to nearest by truncation of a positive number (valid for world coordinates above -200)
		const Vec2 shifted = pos + 200.5;
		return index(Vec2i(shifted) - 200);
	}

	uint32 Grid::index(Vec3 pos) const
	{
		return index(Vec2(pos[0], pos[2]));
	}

	Vec2i Grid::position(uint32 idx) const
	{
		const uint32 width = uint32(resolution[0]);
		return Vec2i(sint32(idx % width), sint32(idx / width)) - gridOffset;
	}

	Vec3 Grid::center(uint32 idx) const
	{
		const Vec2i p = position(idx);
		return Vec3(Real(p[0]), stor(elevations[idx]), Real(p[1]));
	}

	Vec3 Grid::up(uint32 idx) const
	{
		// surface normal from the four neighbouring tile centres (visual use only: path and mana markers)
		const Vec2i p = position(idx);
		const Vec3 west = center(index(p - Vec2i(1, 0)));
		const Vec3 east = center(index(p + Vec2i(1, 0)));
		const Vec3 north = center(index(p - Vec2i(0, 1)));
		const Vec3 south = center(index(p + Vec2i(0, 1)));
		return normalize(cross(south - north, east - west));
	}

	uint32 Grid::neighborDistance(uint32 a, uint32 b) const
	{
		// 24.8-style fixed point of the euclidean distance between the two tile centres, float32 throughout
		return rtos32(distance(center(a), center(b)));
	}

	Holder<Grid> newGrid(Holder<Procedural> procedural)
	{
		Holder<Grid> grid = systemMemory().createHolder<Grid>();
		grid->resolution = Vec2i(MapSize);
		grid->gridOffset = Vec2i(MapOrigin);
		const uint32 total = uint32(MapSize) * uint32(MapSize);

		PointerRangeHolder<const sint16> heights;
		heights.resize(total);
		sint16 *h = const_cast<sint16 *>(heights.data());
		PointerRangeHolder<TileFlags> flags;
		flags.resize(total);
		mtd_mapgen_params params = {};
		params.W = params.H = MapSize;
		params.ox = params.oy = MapOrigin;
		params.radius = MapRadius;
		params.nwp = 0; // newWaypoints() draws the spawn tiles afterwards, from the caller's generator
		if (const b200::SeededProcedural *seeded = dynamic_cast<const b200::SeededProcedural *>(+procedural))
		{
			b200::Device &dev = b200::device();
			std::lock_guard<std::mutex> lock(dev.mutex);
			b200::check(dev, mtd_mapgen_batch(dev.ctx, &params, seeded->s1, seeded->s2, 0, 1, h, (uint16_t *)flags.data(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr));
		}
		else
		{
			for (uint32 i = 0; i < total; i++)
			{
				const Real e = procedural->elevation(Vec2(grid->position(i)));
				h[i] = rtos16(clamp(e, -100, 100));
			}
			b200::Device &dev = b200::device();
			std::lock_guard<std::mutex> lock(dev.mutex);
			b200::check(dev, mtd_grid_from_elevations(dev.ctx, &params, 1, h, (uint16_t *)flags.data()));
		}
		grid->elevations = heights;
		grid->flags = flags;
		return grid;
	}
}
