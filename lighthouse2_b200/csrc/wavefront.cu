// wavefront.cu -- the frame around the two waves: ray generation, minimal diffuse shade, connect, finalize.
// (placeholder until the wavefront stage lands; the raw extend/shadow waves do not depend on it)
#include "rendercore.h"

namespace b200 {

struct WavefrontState { int dummy; };
void DestroyWavefront( WavefrontState* wf ) { delete wf; }

int RenderCore::SettingWavefront( const char* name, float ) { return fail( std::string( "Setting: unknown key " ) + name ); }
int RenderCore::SetTarget( int, int, int ) { return fail( "SetTarget: wavefront stage not built yet" ); }
int RenderCore::SetPointLight( const float*, const float* ) { return fail( "SetLights: wavefront stage not built yet" ); }
int RenderCore::SetMaterials( const void*, int ) { return 0; }
int RenderCore::SetSky( const float* ) { return fail( "SetSkyData: wavefront stage not built yet" ); }
int RenderCore::Render( const void*, int ) { return fail( "Render: wavefront stage not built yet" ); }
int RenderCore::ReadAccumulator( float* ) { return fail( "no accumulator" ); }
int RenderCore::ReadFrame( float* ) { return fail( "no frame" ); }
void* RenderCore::DeviceAccumulator() { return nullptr; }
int RenderCore::SetProbe( int, int ) { return 0; }

} // namespace b200
