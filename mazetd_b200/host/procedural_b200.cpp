// procedural_b200.cpp -- link-time replacement for mazetd's sources/map/procedural.cpp.
//
// The reference's translation unit is compiled INTO this one, unchanged and in place (an #include of the maintainer's own
// tree, nothing is copied): ProceduralImpl -- the RNG draw order of its constructor (procedural.cpp:13-63), the CPU
// elevation (:65-70), material() and sdf() (:79-134, visuals) -- stays exactly the reference's. What changes is
// newProcedural() (:138-141): it remembers the state the thread-local generator had BEFORE the seven draws of the
// constructor, and hands out a forwarding object that carries it. With that state newGrid() (grid_b200.cpp) lets the
// device redo the same draws and sample all 10 201 elevations itself (mtd_mapgen_batch), instead of 10 201 virtual calls
// into the CPU noise; a Procedural that did not come from this function is still sampled through its virtual interface.
#include "mtd_host.h"

#include <cage-core/random.h>

#define newProcedural newProceduralOfTheReference
#include "map/procedural.cpp" // the reference's own file, found through -I <mazetd>/sources
#undef newProcedural

#include "procedural_b200.h"

namespace mazetd
{
	Holder<Procedural> newProcedural()
	{
		const RandomGenerator before = detail::randomGenerator(); // cage/libcore/math/random.cpp:187-194: thread-local
		Holder<b200::SeededProcedural> p = systemMemory().createHolder<b200::SeededProcedural>();
		p->s1 = before.s[0];
		p->s2 = before.s[1];
		p->inner = newProceduralOfTheReference();
		return std::move(p).cast<Procedural>();
	}
}
