// paths_b200.cpp -- link-time replacement for mazetd's sources/map/paths.cpp behind the unchanged sources/grid.h.
//
// Same four definitions, same signatures, same results (bit for bit), the flow-field work done on the GPU through the
// C ABI of libmazetd_b200.so (include/mazetd_b200.h):
//   Directions::update()      reference sources/map/paths.cpp:24-78    -> mtd_grid_upload + mtd_directions_update
//   Waypoints::update()       reference sources/map/paths.cpp:111-139  -> mtd_grid_upload + ONE mtd_waypoints_update
//                                                                         (all layers, all path walks, the aggregates)
//   Waypoints::find()         reference sources/map/paths.cpp:141-169  -> stays on the host: it is a per-call lookup used
//                                                                         all over the game, reading the host mirrors
//   newWaypoints()            reference sources/map/paths.cpp:173-201  -> spawn picking on the host (sequential in the
//                                                                         caller's thread-local RNG stream), then update()
// Build: compile this file and grid_b200.cpp INSTEAD of sources/map/paths.cpp and sources/map/grid.cpp, with
// -I<mazetd>/sources -I<repo>/include, and link libmazetd_b200.so.
#include <vector>

#include "map/generate.h"
#include "mtd_host.h"

#include <cage-core/pointerRangeHolder.h>

namespace mazetd
{
	namespace
	{
		// result arrays are allocated on first use and then reused in place (game code keeps indices into them)
		void ensureArray(Holder<PointerRange<uint32>> &arr, uint32 count)
		{
			if (arr && arr.size() == count)
				return;
			PointerRangeHolder<uint32> fresh;
			fresh.resize(count);
			arr = fresh;
		}

		void uploadGrid(b200::Device &dev, const Grid &grid)
		{
			const int rc = mtd_grid_upload(dev.ctx, grid.resolution[0], grid.resolution[1], grid.gridOffset[0], grid.gridOffset[1], grid.elevations.data(), (const uint16_t *)grid.flags.data());
			b200::check(dev, rc);
		}
	}

	void Directions::update()
	{
		CAGE_ASSERT(grid);
		const uint32 total = numeric_cast<uint32>(grid->flags.size());
		ensureArray(tiles, total);
		ensureArray(distances, total);
		b200::Device &dev = b200::device();
		std::lock_guard<std::mutex> lock(dev.mutex);
		uploadGrid(dev, *grid);
		const uint32 source = tile;
		b200::check(dev, mtd_directions_update(dev.ctx, &source, 1, MTD_M, tiles.data(), distances.data()));
	}

	void Waypoints::update()
	{
		const uint32 count = numeric_cast<uint32>(waypoints.size());
		if (count == 0)
			return;
		const Grid &grid = *waypoints[0]->grid;
		const uint32 total = numeric_cast<uint32>(grid.flags.size());

		std::vector<uint32> targets(count), layerTiles((size_t)count * total), layerDist((size_t)count * total);
		std::vector<uint32> paths((size_t)count * total), pathLength(count), pathDistance(count);
		uint32 aggregates[3] = { 0, 0, m };
		for (uint32 i = 0; i < count; i++)
			targets[i] = waypoints[i]->tile;
		{
			b200::Device &dev = b200::device();
			std::lock_guard<std::mutex> lock(dev.mutex);
			uploadGrid(dev, grid);
			b200::check(dev, mtd_waypoints_update(dev.ctx, targets.data(), (int32_t)count, layerTiles.data(), layerDist.data(), paths.data(), total, pathLength.data(), pathDistance.data(), aggregates));
		}
		for (uint32 i = 0; i < count; i++)
		{
			Waypoint &w = *waypoints[i];
			ensureArray(w.tiles, total);
			ensureArray(w.distances, total);
			detail::memcpy(w.tiles.data(), layerTiles.data() + (size_t)i * total, (size_t)total * sizeof(uint32));
			detail::memcpy(w.distances.data(), layerDist.data() + (size_t)i * total, (size_t)total * sizeof(uint32));
			PointerRangeHolder<uint32> route;
			route.resize(pathLength[i]);
			detail::memcpy(route.data(), paths.data() + (size_t)i * total, (size_t)pathLength[i] * sizeof(uint32));
			w.fullPath = route;
			w.fullDistance = pathDistance[i];
		}
		avgFullDistance = aggregates[0];
		minFullDistance = aggregates[1];
		minDistanceSpawner = aggregates[2];
	}

	namespace
	{
		// Depth-first enumeration of the orders in which the remaining waypoints can be visited, candidates taken in
		// ascending index order at every depth: that IS lexicographic permutation order, so "first strictly smaller cost
		// wins" picks the same order as the reference's next_permutation loop. Costs are uint32 sums that wrap, as there.
		struct RouteSearch
		{
			const Waypoints &wps;
			uint32 remaining[32];
			uint32 remainingCount = 0;
			uint32 bestCost = m;
			uint32 bestFirst = m;

			explicit RouteSearch(const Waypoints &wps) : wps(wps) {}

			void descend(uint32 usedMask, uint32 depth, uint32 position, uint32 costSoFar, uint32 first)
			{
				if (depth == remainingCount)
				{
					if (costSoFar < bestCost)
					{
						bestCost = costSoFar;
						bestFirst = first;
					}
					return;
				}
				for (uint32 k = 0; k < remainingCount; k++)
				{
					if (usedMask & (1u << k))
						continue;
					const uint32 w = remaining[k];
					const uint32 leg = wps.waypoints[w]->distances[position];
					descend(usedMask | (1u << k), depth + 1, wps.waypoints[w]->tile, costSoFar + leg, depth == 0 ? w : first);
				}
			}
		};
	}

	Waypoints::FindResult Waypoints::find(uint32 currentPosition, uint32 visitedWaypointsBits) const
	{
		RouteSearch search(*this);
		const uint32 count = numeric_cast<uint32>(waypoints.size());
		for (uint32 i = 0; i < count; i++)
			if (!(visitedWaypointsBits & (1u << i)))
				search.remaining[search.remainingCount++] = i;
		FindResult result;
		if (search.remainingCount == 0)
			return result;
		search.descend(0, 0, currentPosition, 0, m);
		// like the reference, the search starts from cost m and only a strictly smaller cost is accepted: when every order
		// sums to m (or wraps above it) the answer stays { m, m }
		if (search.bestFirst == m)
			return result;
		result.tile = waypoints[search.bestFirst]->tiles[currentPosition];
		result.distance = search.bestCost;
		return result;
	}

	Holder<Waypoints> newWaypoints(Holder<Grid> grid)
	{
		constexpr uint32 spawnerCount = 4;
		Holder<Waypoints> result = systemMemory().createHolder<Waypoints>();
		PointerRangeHolder<Holder<Waypoints::Waypoint>> list;
		list.reserve(spawnerCount);
		const uint32 total = numeric_cast<uint32>(grid->flags.size());
		while (list.size() < spawnerCount)
		{
			// rejection sampling in the caller's RNG stream: one randomRange draw per attempt, like the reference
			const uint32 candidate = randomRange(0u, total);
			if (any(grid->flags[candidate] & (TileFlags::Invalid | TileFlags::Waypoint)))
				continue;
			grid->flags[candidate] |= TileFlags::Waypoint;
			Holder<Waypoints::Waypoint> w = systemMemory().createHolder<Waypoints::Waypoint>();
			w->grid = grid.share();
			w->tile = candidate;
			list.push_back(std::move(w));
		}
		result->waypoints = list;
		result->update();
		return result;
	}
}
