// procedural_b200.h -- the forwarding Procedural handed out by the replacement newProcedural() (procedural_b200.cpp)
#ifndef mazetd_b200_procedural_h
#define mazetd_b200_procedural_h

#include "map/generate.h"

namespace mazetd
{
	namespace b200
	{
		struct SeededProcedural : public Procedural
		{
			Holder<Procedural> inner; // the reference's ProceduralImpl
			uint64 s1 = 0, s2 = 0;    // generator state before its constructor drew from it: the device repeats the draws

			Real elevation(const Vec2 &pos) override { return inner->elevation(pos); }
			Real sdf(const Vec3 &pos) override { return inner->sdf(pos); }
			void material(const Vec3 &position, TileFlags flags, Vec3 &albedo, Real &roughness) override { inner->material(position, flags, albedo, roughness); }
		};
	}
}

#endif
