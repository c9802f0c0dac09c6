// b2fwn_mc.cu -- marching cubes on device slabs (placeholder until the kernels land; fails loudly).
#include "b2fwn_internal.h"
namespace b2
{
void freeMc( McResult& r )
{
    cudaFree( r.dVerts ); cudaFree( r.dTris );
    r = McResult{};
}
int marchingCubesDevice( const float*, const int32_t*, const b2fwn_mc_params&, McResult&, cudaStream_t, uint64_t* )
{
    setError( "marching cubes: not built yet" );
    return B2FWN_ERR_INVALID;
}
} // namespace b2
