// Batch of independent calibrations: one thread block per problem (filled in below).
#pragma once
