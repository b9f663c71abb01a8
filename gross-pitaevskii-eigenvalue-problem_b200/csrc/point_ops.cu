// placeholder until the loss kernels land
namespace gpe { }
