// mtd_host.h -- glue shared by the replacement bodies in this directory: one process-wide mtd_ctx behind a mutex, and the
// translation of C-ABI status codes into cage exceptions (the reference reports failures as C++ exceptions through
// CAGE_THROW_ERROR, externals/cage/sources/include/cage-core/core.h:37-64; no exception crosses the C boundary itself).
#ifndef mazetd_b200_host_h
#define mazetd_b200_host_h

#include <cstdio>
#include <cstdlib>
#include <mutex>

#include <cage-core/core.h>

#include "mazetd_b200.h"

namespace mazetd
{
	namespace b200
	{
		// The game drives the map from the control thread and from one task-pool thread (map generation); every adapter
		// call takes this lock for its whole upload + compute + download sequence, so results never interleave.
		struct Device
		{
			std::mutex mutex;
			mtd_ctx *ctx = nullptr;

			Device()
			{
				// MAZETD_B200_DEVICE selects the CUDA ordinal (default 0). There is no CPU fallback: without a device the
				// first map operation throws, exactly like any other unrecoverable engine error.
				int ordinal = 0;
				if (const char *e = std::getenv("MAZETD_B200_DEVICE"))
					ordinal = std::atoi(e);
				if (mtd_create(&ctx, ordinal, nullptr) != MTD_OK)
				{
					ctx = nullptr;
					CAGE_THROW_ERROR(cage::Exception, "mazetd_b200: no usable CUDA device (the GPU map path has no CPU fallback)");
				}
			}

			~Device()
			{
				if (ctx)
					mtd_destroy(ctx);
			}
		};

		inline Device &device()
		{
			static Device d;
			return d;
		}

		inline void check(Device &d, int rc)
		{
			if (rc == MTD_OK)
				return;
			std::fprintf(stderr, "mazetd_b200: %s (code %d)\n", mtd_last_error(d.ctx), rc);
			CAGE_THROW_ERROR(cage::Exception, "mazetd_b200: device map operation failed");
		}
	}
}

#endif
