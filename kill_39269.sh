ssh test@127.0.0.1 "kill -9 2516"
ssh test@127.0.0.1 "kill -9 2517"
ssh test@127.0.0.1 "kill -9 2518"
kill -9 2109
