"""Writes tests/golden/flow_params_*.json with the REFERENCE's own
FlowParamsEncoder (flow/utils/rllib.py:22-60) from flow_params built with the
reference's classes — the file format RLlib experiments store. Run in the build
container (needs /root/reference)."""
import json
import os
import types

from . import ref_shim

ref_shim.install()

import flow.controllers as rc  # noqa: E402
import flow.core.params as rp  # noqa: E402
import flow.envs as re_  # noqa: E402
import flow.envs.multiagent as rm  # noqa: E402
import flow.networks as rn  # noqa: E402
from flow.utils.rllib import FlowParamsEncoder  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def ref_api():
    return types.SimpleNamespace(
        VehicleParams=rp.VehicleParams, SumoCarFollowingParams=rp.SumoCarFollowingParams,
        SumoParams=rp.SumoParams, EnvParams=rp.EnvParams, NetParams=rp.NetParams,
        InitialConfig=rp.InitialConfig, InFlows=rp.InFlows, IDMController=rc.IDMController,
        RLController=rc.RLController, ContinuousRouter=rc.ContinuousRouter,
        WaveAttenuationPOEnv=re_.WaveAttenuationPOEnv, RingNetwork=rn.RingNetwork,
        AccelEnv=re_.AccelEnv, HighwayNetwork=rn.HighwayNetwork,
        MultiAgentWaveAttenuationPOEnv=rm.MultiAgentWaveAttenuationPOEnv,
        MergePOEnv=re_.MergePOEnv, MergeNetwork=rn.MergeNetwork,
        FigureEightNetwork=rn.FigureEightNetwork,
        MERGE_NET_PARAMS=__import__('flow.networks.merge', fromlist=['x']).ADDITIONAL_NET_PARAMS,
        FIGURE_EIGHT_NET_PARAMS=__import__('flow.networks.figure_eight',
                                           fromlist=['x']).ADDITIONAL_NET_PARAMS)


def main():
    from tests.test_flow_api import ring_flow_params, c1_flow_params
    from tests.test_rllib_compat import highway_flow_params
    from tests.test_gpu_merge import merge_flow_params
    from tests.test_gpu_figure8 import figure8_flow_params
    api = ref_api()
    for name, fp in (('c2', ring_flow_params(api, noise=0.2)), ('c1', c1_flow_params(api)),
                     ('highway', highway_flow_params(api)), ('merge', merge_flow_params(api)),
                     ('figure8', figure8_flow_params(api, noise=0.2))):
        txt = json.dumps(fp, cls=FlowParamsEncoder, sort_keys=True, indent=4)
        open(os.path.join(HERE, 'flow_params_%s.json' % name), 'w').write(txt)
        print(name, len(txt))


if __name__ == '__main__':
    main()
