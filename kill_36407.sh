ssh test@127.0.0.1 "kill -9 2997"
ssh test@127.0.0.1 "kill -9 2998"
ssh test@127.0.0.1 "kill -9 2999"
kill -9 2946
