ssh test@127.0.0.1 "kill -9 2911"
ssh test@127.0.0.1 "kill -9 2912"
kill -9 2109
