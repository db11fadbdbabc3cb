# parameters.h:16 -- the one-token fix g++ needs (SURVEY.md D8)
s/^#define STCONST static const[[:space:]]*$/#define STCONST static constexpr/
# parameters.h:77 -- let a config switch the intracellular-throw flag off
s/^([[:space:]]*)#define AVOID_INTRACELLULAR_THROW[[:space:]]*$/\1#ifndef REF_NO_AVOID_INTRACELLULAR_THROW\n\1#define AVOID_INTRACELLULAR_THROW\n\1#endif/
# every knob a BASELINE.json config varies: keep the shipped initialiser as the default
s/^([[:space:]]*STCONST[[:space:]]+(int|double)[[:space:]]+)(num_water|num_cells|cell_r|mmoment|bound|water_start_bound|hashDim|maxNeighbors|border|scale|sprintSteps|totaltime|num_threads)([[:space:]]*=[[:space:]]*)([^;]+);(.*)$/#ifdef REF_\3\n\1\3\4REF_\3;\6\n#else\n\1\3\4\5;\6\n#endif/
