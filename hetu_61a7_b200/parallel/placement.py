def assign_placement(config, nodes):
    pass
def split_stages(node_list, stage_groups, boundaries=None):
    pass
